"""CPU: the C-ABI library loads, exports every symbol include/pn_b200.h declares, and the ctypes
signature table covers the header one to one.  No compute calls (no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pn_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pn_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def built_lib():
    from pacednegatives_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _lib


def test_header_symbols_exported(built_lib):
    names = declared_symbols()
    assert len(names) >= 15
    raw = ctypes.CDLL(built_lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"{n} declared in pn_b200.h but not exported"


def test_signature_table_matches_header(built_lib):
    assert sorted(built_lib.SIGNATURES) == declared_symbols()
    lib = built_lib.lib()
    assert lib.pn_version() == 100
    assert lib.pn_launch_count() >= 0


def test_no_cpu_fallback(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        built_lib.require_gpu()
    import pacednegatives_b200.weights as w
    with pytest.raises(RuntimeError):
        w.LCEWeights(0.5)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "pacednegatives_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f


def test_checkpoint_directory_loader(tmp_path):
    """load_checkpoint reads what HF save_pretrained writes (lceT5.py:106 / eval.py:14 wire format)."""
    import torch
    from transformers import T5Config, T5ForConditionalGeneration
    from pacednegatives_b200.scorer import dims_from_hf_config, hf_config, load_checkpoint, T5Dims, _tensor_shape
    cfg = T5Config(vocab_size=96, d_model=64, d_kv=16, d_ff=128, num_layers=2, num_decoder_layers=1, num_heads=4,
                   dropout_rate=0.0, decoder_start_token_id=0, pad_token_id=0, eos_token_id=1,
                   feed_forward_proj="relu", tie_word_embeddings=True)
    hf = T5ForConditionalGeneration(cfg)
    hf.save_pretrained(str(tmp_path))
    dims, sd = load_checkpoint(str(tmp_path))
    assert dims == T5Dims(vocab=96, d_model=64, d_kv=16, d_ff=128, n_layers=2, n_dec_layers=1, n_heads=4)
    ref = hf.state_dict()
    assert torch.equal(sd["shared.weight"], ref["shared.weight"])
    for k, v in sd.items():
        if k not in ("encoder.embed_tokens.weight", "decoder.embed_tokens.weight", "lm_head.weight"):
            assert _tensor_shape(k, v.numel(), dims) == tuple(v.shape), k
    back = T5Config(**hf_config(dims))
    assert dims_from_hf_config(back.to_dict()) == dims
    import pytest
    with pytest.raises(ValueError):
        dims_from_hf_config({**hf_config(dims), "feed_forward_proj": "gated-gelu"})


def test_host_pacing_helpers_without_gpu():
    """interpolate_scalar (old/interp.py:10-18) and ChangeDifficulty (lceT5.py:77-80) are pure host logic."""
    import types
    from pacednegatives_b200.wrappers import interpolate_scalar
    from pacednegatives_b200.lceT5 import ChangeDifficulty, linear_schedule_with_warmup
    f = interpolate_scalar(0.0, 1.0, 24)
    assert [f(s) for s in (0, 4, 24, 25, 1000)] == [0.0, 4 * (1.0 / 24), 24 * (1.0 / 24), 1.0, 1.0]
    ds = types.SimpleNamespace(weight=0.0)
    trainer = types.SimpleNamespace(train_dataloader=types.SimpleNamespace(dataset=ds))
    cb = ChangeDifficulty()
    cb.on_train_batch_start(trainer, types.SimpleNamespace(difficulty=0.4), None, 0)
    assert ds.weight == 0.4
    cb.on_train_batch_start(trainer, types.SimpleNamespace(difficulty=3.0), None, 1)
    assert ds.weight == 1 - 1e-10
    assert linear_schedule_with_warmup(1, 2, 12) == 0.5 and linear_schedule_with_warmup(7, 2, 12) == 0.5

"""Row b on the GPU: the reference's training driver (train/train_lce.py:20-103, restated call for call in
tests/helpers/train_lce_driver.py and compared with the reference file by tests/test_dropin_cpu.py) runs against
pacednegatives_b200 with `pacednegatives` aliased to it: LCEModel(hparams) with model_name -> a local checkpoint
directory (config.json + model.safetensors + a SentencePiece model trained here offline), model.tokenizer feeding
LCEDataModule, ChangeDifficulty, trainer.fit (a stub of Lightning's manual-optimisation loop), save_pretrained."""
import importlib
import json
import os
import random
import sys
import types

import numpy as np
import pytest
import torch

from oracle import t5 as ot

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORDS = [f"w{i}" for i in range(300)] + ["Query:", "Document:", "Relevant:", "true", "false", "the", "of", "and"]


class Corpus:
    def __init__(self, n_docs, n_queries, seed=0):
        rnd = random.Random(seed)
        self.docs = [" ".join(rnd.choice(WORDS[:300]) for _ in range(rnd.randint(6, 18))) for _ in range(n_docs)]
        self.queries = [" ".join(rnd.choice(WORDS[:300]) for _ in range(rnd.randint(2, 5))) for _ in range(n_queries)]

    def docs_iter(self):
        for i, t in enumerate(self.docs):
            yield {"doc_id": str(i), "text": t}

    def queries_iter(self):
        for i, t in enumerate(self.queries):
            yield {"query_id": str(i), "text": t}


def make_checkpoint(path, seed=41, dropout_rate=0.1):
    """A local 't5-*'-style directory: SentencePiece vocabulary (284 pieces + 100 extra ids = 384), config.json with
    the checkpoint's dropout_rate, and weights that carry the legacy / tied keys stock checkpoints do."""
    import sentencepiece as spm
    from safetensors.torch import save_file
    os.makedirs(path, exist_ok=True)
    rnd = random.Random(1)
    with open(os.path.join(path, "corpus.txt"), "w") as f:
        for _ in range(3000):
            f.write(" ".join(rnd.choice(WORDS) for _ in range(12)) + "\n")
    spm.SentencePieceTrainer.train(input=os.path.join(path, "corpus.txt"), model_prefix=os.path.join(path, "spiece"),
                                   vocab_size=284, model_type="unigram", pad_id=0, eos_id=1, unk_id=2, bos_id=-1,
                                   hard_vocab_limit=False, minloglevel=2)
    dims = ot.T5Dims(vocab=384, d_model=128, d_kv=64, d_ff=256, n_layers=2, n_dec_layers=2, n_heads=2)
    w = ot.init_weights(dims, seed=seed)
    sd = {k: torch.as_tensor(v).contiguous() for k, v in w.items()}
    # what a stock t5 checkpoint also holds: HF ignores the first on load, the other three are ties
    sd["decoder.block.0.layer.1.EncDecAttention.relative_attention_bias.weight"] = torch.zeros(32, dims.n_heads)
    for alias in ("encoder.embed_tokens.weight", "decoder.embed_tokens.weight", "lm_head.weight"):
        sd[alias] = sd["shared.weight"].clone()
    save_file(sd, os.path.join(path, "model.safetensors"), metadata={"format": "pt"})
    cfg = {"architectures": ["T5ForConditionalGeneration"], "model_type": "t5", "vocab_size": dims.vocab,
           "d_model": dims.d_model, "d_kv": dims.d_kv, "d_ff": dims.d_ff, "num_layers": 2, "num_decoder_layers": 2,
           "num_heads": dims.n_heads, "relative_attention_num_buckets": 32, "relative_attention_max_distance": 128,
           "layer_norm_epsilon": 1e-6, "dropout_rate": dropout_rate, "feed_forward_proj": "relu",
           "is_encoder_decoder": True, "initializer_factor": 1.0, "tie_word_embeddings": True, "pad_token_id": 0,
           "eos_token_id": 1, "decoder_start_token_id": 0, "n_positions": 512,
           "task_specific_params": {"summarization": {"max_length": 200}}}
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(cfg, f)
    return dims, w


@pytest.fixture
def aliased(monkeypatch):
    """pacednegatives -> pacednegatives_b200, the lightning stub, and an ir_datasets whose load() returns our corpus"""
    sys.path.insert(0, os.path.join(ROOT, "tests", "helpers"))
    import lightning_stub
    import pacednegatives_b200
    import pacednegatives_b200.lceT5
    saved = {k: sys.modules.get(k) for k in ("pacednegatives", "pacednegatives.lceT5", "lightning",
                                             "lightning.pytorch", "ir_datasets", "train_lce_driver")}
    lightning_stub.install()
    sys.modules["pacednegatives"] = pacednegatives_b200
    sys.modules["pacednegatives.lceT5"] = pacednegatives_b200.lceT5
    corpus = Corpus(64, 12)
    irds = types.ModuleType("ir_datasets")
    irds.load = lambda name: corpus
    sys.modules["ir_datasets"] = irds
    sys.modules.pop("train_lce_driver", None)
    driver = importlib.import_module("train_lce_driver")
    yield driver, corpus
    for k, v in saved.items():
        if v is None:
            sys.modules.pop(k, None)
        else:
            sys.modules[k] = v
    sys.path.remove(os.path.join(ROOT, "tests", "helpers"))


def test_reference_driver_runs_against_the_drop_in(tmp_path, aliased):
    driver, corpus = aliased
    ckpt = str(tmp_path / "t5-tiny")
    dims, w0 = make_checkpoint(ckpt)
    rng = np.random.RandomState(5)
    C = 40
    rows = [{"query_id": str(q), "doc_id_a": str(int(rng.randint(64))),
             "doc_id_b": [str(int(x)) for x in rng.permutation(64)[:C]]} for q in range(12)]
    data = str(tmp_path / "train.json")
    with open(data, "w") as f:
        json.dump(rows, f)
    out_dir = str(tmp_path / "out")
    np.random.seed(7)
    with pytest.warns(UserWarning, match="ignored 1 checkpoint tensor"):
        model, trainer = driver.main(data, "synthetic", out_dir, total_steps=64, eta=20.0, batch_size=4, lr=1e-3,
                                     meta_lr=1e-2, var=0.01, n=3, warmup_steps=8, use_mean=True, num_workers=0,
                                     max_steps=3, model_name=ckpt)
    # the reference's attribute surface
    from pacednegatives_b200.lceT5 import LCEModel, TokenCELoss
    assert isinstance(model, LCEModel) and model.tokenizer is not None and isinstance(model.loss_fn, TokenCELoss)
    assert model.automatic_optimization is False and model.hparams.model_name == ckpt
    assert model.dropout_rate == 0.1            # the checkpoint's config.json value: the reference's train-mode regime
    assert model.hparams.total_steps == 16 and model.hparams.warmup_steps == 2
    assert len(trainer.outputs) == 3
    for o in trainer.outputs:
        assert list(o.keys()) == ["main_loss", "meta_loss", "progress_bar", "log"]
        assert torch.isfinite(o["main_loss"]) and torch.isfinite(o["meta_loss"])
    # two optimisers + two schedulers, stepped three times: linear warm-up over 2 steps then linear decay to 16
    (meta_opt, opt), (meta_sched, sched) = model.optimizers(), model.lr_schedulers()
    assert abs(sched.get_last_lr()[0] - 1e-3 * (16 - 3) / (16 - 2)) < 1e-12
    assert abs(meta_sched.get_last_lr()[0] - 1e-2 * (16 - 3) / (16 - 2)) < 1e-12
    assert model.weights.eta.item() != 20.0 and model.difficulty == model.weights.eta.item()
    # ChangeDifficulty pushed eta (clipped) into the dataset before each batch
    assert trainer.train_dataloader.dataset.weight == 1 - 1e-10     # eta ~ 20 -> clipped
    # save_pretrained: loads in HF, moved away from the initial weights, keeps the checkpoint's config fields
    final = os.path.join(out_dir, "finalmodel")
    cfg = json.load(open(os.path.join(final, "config.json")))
    assert cfg["dropout_rate"] == 0.1 and cfg["n_positions"] == 512 and "task_specific_params" in cfg
    from transformers import T5ForConditionalGeneration
    hf = T5ForConditionalGeneration.from_pretrained(final)
    moved = (hf.state_dict()["encoder.block.0.layer.0.SelfAttention.q.weight"]
             - torch.as_tensor(w0["encoder.block.0.layer.0.SelfAttention.q.weight"])).abs().max().item()
    assert 1e-4 < moved < 1e-2      # three AdamW steps at lr <= 1e-3


def test_first_step_loss_matches_hf_on_the_tokenised_batch(tmp_path, aliased):
    """The same collate -> training_step path against HF fp32 on the very batch the data module produced (padded to
    the longest prompt, no attention mask: lceT5.py:53 drops it, so pad tokens are attended here and there alike)."""
    driver, corpus = aliased
    ckpt = str(tmp_path / "t5-tiny")
    dims, w0 = make_checkpoint(ckpt)
    from pacednegatives_b200.lceT5 import LCEModel, LCEDataModule
    model = LCEModel({"model_name": ckpt, "total_steps": 10, "eta": 20.0, "batch_size": 4, "lr": 1e-3,
                      "meta_lr": 1e-2, "n": 3, "warmup_steps": 0, "use_mean": True, "ignore_index": -100,
                      "dropout_rate": 0.0})     # HF below runs in eval mode; the default is the checkpoint's 0.1
    assert model.dropout_rate == 0.0
    rng = np.random.RandomState(5)
    rows = [{"query_id": str(q), "doc_id_a": str(int(rng.randint(64))),
             "doc_id_b": [str(int(x)) for x in rng.permutation(64)[:40]]} for q in range(4)]
    data = str(tmp_path / "train.json")
    with open(data, "w") as f:
        json.dump(rows, f)
    dm = LCEDataModule(data, "synthetic", model.tokenizer, 4, False, False, var=0.01, n=3,
                       init_weight=0.5, num_workers=0)
    dm.setup()
    np.random.seed(3)
    p, n = next(iter(dm.train_dataloader()))
    assert p["input_ids"].shape[0] == 4 and n["input_ids"].shape[0] == 12 and p["labels"].shape[1] == 2
    out = model.training_step((p, n), 0)
    from transformers import T5Config, T5ForConditionalGeneration
    hf = T5ForConditionalGeneration.from_pretrained(ckpt).eval()
    with torch.no_grad():
        pl = hf(input_ids=p["input_ids"], labels=p["labels"]).logits
        nl = hf(input_ids=n["input_ids"], labels=n["labels"]).logits
    ce = torch.nn.CrossEntropyLoss(ignore_index=-100, reduction="none")
    pce = ce(pl.view(-1, pl.size(-1)), p["labels"].view(-1))
    nce = ce(nl.view(-1, nl.size(-1)), n["labels"].view(-1)).view(-1, 3).mean(1)
    ref = float((pce + nce).mean())
    assert abs(float(out["main_loss"].detach()) - ref) < 2e-2, (float(out["main_loss"]), ref)

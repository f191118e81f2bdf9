"""NewWrapper (pairwrapper.py:208-283) on the B200 kernels vs the golden vectors the reference's own
meta_loop / main_loop produced on CPU (tests/golden/make_golden.py::gold_newwrapper), and checkpoint I/O."""
import os
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _tiny(seed):
    from oracle import t5 as ot
    from pacednegatives_b200.scorer import T5Dims
    dims = ot.T5Dims(vocab=384, d_model=128, d_kv=64, d_ff=256, n_layers=2, n_dec_layers=2, n_heads=2)
    w = ot.init_weights(dims, seed=seed)
    return ot, dims, T5Dims(**dims.__dict__), w


def test_newwrapper_vs_reference_golden():
    from pacednegatives_b200.lceT5 import MonoT5Model
    from pacednegatives_b200.pairwrapper import NewWrapper, _Schedule
    from pacednegatives_b200.weights import Weights
    g = np.load(os.path.join(HERE, "golden", "newwrapper_golden.npz"))
    ot, odims, dims, w = _tiny(int(g["seed"]))
    B, lr, warm, total, rel_id, nrel_id = [float(x) for x in g["hp"]]
    B, warm, total, rel_id, nrel_id = int(B), int(warm), int(total), int(rel_id), int(nrel_id)
    pids, nids = g["pids"], g["nids"]
    nb = pids.shape[0]

    class Tok:
        def __call__(self, x, padding=True, return_tensors="pt"):
            if isinstance(x, tuple):
                ids = torch.tensor((pids if x[0] == "p" else nids)[x[1]])
            else:
                ids = torch.tensor([[rel_id if t == "true" else nrel_id, 1] for t in x])
            return types.SimpleNamespace(input_ids=ids)

    seen = []

    class Loader:
        def get_batch(self, j, weights):
            seen.append(weights.detach().cpu().numpy().copy())
            return ("p", j), ("n", j), ["true"] * B, ["false"] * B

        def __len__(self):
            return nb * B

    sd = {k: torch.as_tensor(v) for k, v in w.items()}
    wrap = NewWrapper("synthetic", "tiny", B, lambda: MonoT5Model(dims, sd), Tok(), lr, -100)
    before = wrap.model.parameters_flat().clone()
    wrap.weights = Weights((nb, B))
    wrap.train_loader = Loader()
    wrap.scheduler = _Schedule(lr, warm, total)
    for j in range(nb):
        wrap.meta_loop(j)
        row = wrap.weights.weight[j].detach().cpu().numpy()
        # sigmoid(1 - lr (ce/B - 1)): a 2e-2 error of a bf16 loss moves a row by < 2e-6
        assert np.abs(row - g[f"row{j}"]).max() < 1e-5, (j, row, g[f"row{j}"])
        loss = wrap.main_loop(j)
        assert abs(loss - float(g[f"loss{j}"])) < 5e-2 * max(1.0, float(g[f"loss{j}"]) / 5), (j, loss, g[f"loss{j}"])
        assert abs(wrap.scheduler.get_last_lr()[0] - float(g[f"lr{j}"])) < 1e-9
    assert np.abs(np.stack(seen) - g["seen"]).max() < 1e-5
    after = wrap.model.parameters_flat()
    for name, off, numel in wrap.model.named_offsets():
        ref = float(g["dnorm/" + name])
        got = float((after[off:off + numel] - before[off:off + numel]).norm())
        assert abs(got - ref) <= 0.05 * ref + 1e-6, (name, got, ref)


def test_stdwrapper_lookahead_vs_reference_golden():
    """StdWrapper (pairwrapper.py:91-206): the look-ahead model theta' = theta - lr grad (embedding excluded, as
    transformers 5.x makes the reference behave), the per-sample CE under theta and theta', the table rows, the
    main-loop losses (negatives through theta') and the parameter movement, against the reference's own
    meta_loop / main_loop (tests/golden/make_golden.py::gold_stdwrapper)."""
    from pacednegatives_b200.lceT5 import MonoT5Model
    from pacednegatives_b200.pairwrapper import StdWrapper, _Schedule
    from pacednegatives_b200.weights import Weights
    g = np.load(os.path.join(HERE, "golden", "stdwrapper_golden.npz"))
    ot, odims, dims, w = _tiny(int(g["seed"]))
    B, lr, warm, total, rel_id, nrel_id = [float(x) for x in g["hp"]]
    B, warm, total, rel_id, nrel_id = int(B), int(warm), int(total), int(rel_id), int(nrel_id)
    pids, nids = g["pids"], g["nids"]
    nb = pids.shape[0]

    class Tok:
        def __call__(self, x, padding=True, return_tensors="pt"):
            if isinstance(x, tuple):
                ids = torch.tensor((pids if x[0] == "p" else nids)[x[1]])
            else:
                ids = torch.tensor([[rel_id if t == "true" else nrel_id, 1] for t in x])
            return types.SimpleNamespace(input_ids=ids)

    seen = []

    class Loader:
        def get_batch(self, j, weights):
            seen.append(weights.detach().cpu().numpy().copy())
            return ("p", j), ("n", j), ["true"] * B, ["false"] * B

        def __len__(self):
            return nb * B

    sd = {k: torch.as_tensor(v) for k, v in w.items()}
    wrap = StdWrapper("synthetic", "tiny", B, lambda: MonoT5Model(dims, sd), Tok(), lr, -100)
    assert wrap.meta_model is not wrap.model
    ces = []
    inner = wrap.loss_fn
    wrap.loss_fn = lambda logits, target: (lambda o: (ces.append(o.detach().cpu().numpy().copy()), o)[1])(
        inner(logits, target))
    before = wrap.model.parameters_flat().clone()
    wrap.weights = Weights((nb, B))
    wrap.weights.weight.data[:] = torch.tensor(g["w0"]).cuda()
    wrap.train_loader = Loader()
    wrap.scheduler = _Schedule(lr, warm, total)
    for j in range(nb):
        del ces[:]
        wrap.meta_loop(j)
        ref = g[f"meta_ce{j}"]                     # ce_p(theta), ce_n(theta), ce_p(theta'), ce_n(theta')
        assert np.abs(ces[0] - ref[0]).max() < 3e-2 and np.abs(ces[1] - ref[1]).max() < 3e-2, (j, ces[:2], ref[:2])
        # the look-ahead moved the negatives' CE by up to ~2.8 in the golden: theta' must reproduce it
        assert np.abs(ces[-1] - ref[3]).max() < 6e-2, (j, ces[-1], ref[3])
        row = wrap.weights.weight[j].detach().cpu().numpy()
        assert np.abs(row - g[f"row{j}"]).max() < 1e-3, (j, row, g[f"row{j}"])     # lr/B * 6e-2 * sigmoid' < 1e-4
        del ces[:]
        loss = wrap.main_loop(j)
        assert abs(loss - float(g[f"loss{j}"])) < 5e-2, (j, loss, g[f"loss{j}"])
        assert np.abs(ces[1] - g[f"main_ce{j}"][1]).max() < 6e-2          # negatives scored by the look-ahead model
        assert abs(wrap.scheduler.get_last_lr()[0] - float(g[f"lr{j}"])) < 1e-8
    assert np.abs(np.stack(seen) - g["seen"]).max() < 1e-3
    after = wrap.model.parameters_flat()
    for name, off, numel in wrap.model.named_offsets():
        ref = float(g["dnorm/" + name])
        got = float((after[off:off + numel] - before[off:off + numel]).norm())
        assert abs(got - ref) <= 0.05 * ref + 1e-6, (name, got, ref)
    # the look-ahead may also include the embedding (older transformers tie by module): runs, rows stay close (the
    # embedding's share of the step is small at this size)
    wrap2 = StdWrapper("synthetic", "tiny", B, lambda: MonoT5Model(dims, sd), Tok(), lr, -100,
                       lookahead_updates_embedding=True)
    wrap2.weights = Weights((nb, B))
    wrap2.weights.weight.data[:] = torch.tensor(g["w0"]).cuda()
    wrap2.train_loader, wrap2.scheduler = Loader(), _Schedule(lr, warm, total)
    wrap2.meta_loop(0)
    assert np.abs(wrap2.weights.weight[0].detach().cpu().numpy() - g["row0"]).max() < 1e-2


@pytest.mark.parametrize("safe", [True, False])
def test_save_pretrained_round_trip(tmp_path, safe):
    """save_pretrained -> HF T5ForConditionalGeneration.from_pretrained (train/train_lce.py:103, eval.py:14) and
    back through MonoT5Scorer.from_pretrained: same tensors, same scores."""
    from transformers import T5ForConditionalGeneration
    from pacednegatives_b200.scorer import MonoT5Scorer
    ot, odims, dims, w = _tiny(11)
    sd = {k: torch.as_tensor(v) for k, v in w.items()}
    m = MonoT5Scorer(dims, sd, rel_id=117, nrel_id=213)
    out = str(tmp_path / "ckpt")
    m.save_pretrained(out, safe_serialization=safe)
    hf = T5ForConditionalGeneration.from_pretrained(out).eval()
    hsd = hf.state_dict()
    for k, v in w.items():
        assert np.array_equal(hsd[k].numpy(), v), k
    assert hf.config.feed_forward_proj == "relu" and hf.config.tie_word_embeddings
    m2 = MonoT5Scorer.from_pretrained(out, rel_id=117, nrel_id=213)
    rng = np.random.RandomState(5)
    ids = rng.randint(2, odims.vocab, size=(6, 128)).astype(np.int64)
    ids[:, -1] = 1
    assert torch.equal(m.score(ids), m2.score(ids))
    with torch.no_grad():
        dec = torch.zeros(6, 1, dtype=torch.long)
        z = hf(input_ids=torch.tensor(ids), decoder_input_ids=dec).logits[:, 0, [117, 213]]
        ref = torch.log_softmax(z, dim=-1)[:, 0]
    assert (m.score(ids).cpu() - ref).abs().max() < 2e-2

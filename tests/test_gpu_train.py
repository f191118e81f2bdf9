"""GPU parity of the training step: LCEModel.training_step on the B200 kernels vs the golden
vectors written by the REFERENCE's own training_step (tests/golden/make_golden.py::gold_train: unmodified
lceT5.LCEModel.training_step driving HF T5ForConditionalGeneration in fp32 with transformers.AdamW's update rule).
Tolerances: north_star's 2e-2 absolute on losses (bf16 tensor-core arithmetic against fp32), each gradient tensor's
norm within 3 % and its leading entries within 10 % of their scale (bf16 operands in every GEMM of the backward)."""
import os

import numpy as np
import pytest
import torch

from oracle import t5 as ot

pytestmark = pytest.mark.gpu


def load_case(golden_dir):
    g = np.load(os.path.join(golden_dir, "train_golden.npz"))
    v, d, kv, ff, L, Ld, H = [int(x) for x in g["dims"]]
    dims = ot.T5Dims(vocab=v, d_model=d, d_kv=kv, d_ff=ff, n_layers=L, n_dec_layers=Ld, n_heads=H)
    n, use_mean, lr, meta_lr, warm, total, eta = [float(x) for x in g["hp"]]
    hp = dict(n=int(n), use_mean=bool(use_mean), ignore_index=-100, lr=lr, meta_lr=meta_lr, warmup_steps=int(warm),
              total_steps=int(total), eta=eta)
    batch = ({"input_ids": torch.tensor(g["pids"]), "labels": torch.tensor(g["plab"])},
             {"input_ids": torch.tensor(g["nids"]), "labels": torch.tensor(g["nlab"])})
    return g, dims, hp, batch


@pytest.mark.parametrize("elide,xattn", [(True, True), (False, True), (True, False)])
def test_training_step_vs_reference_golden(golden_dir, elide, xattn):
    from pacednegatives_b200.lceT5 import LCEModel
    from pacednegatives_b200.scorer import T5Dims
    g, dims, hp, batch = load_case(golden_dir)
    w = ot.init_weights(dims, seed=int(g["seed"]))
    m = LCEModel(hp, dims=T5Dims(**dims.__dict__), state_dict={k: torch.as_tensor(v) for k, v in w.items()},
                 elide_no_grad_forward=elide, xattn=xattn)
    out = m.training_step(batch, 0)
    assert abs(float(out["main_loss"].detach()) - float(g["main_loss0"])) < 2e-2
    assert abs(float(out["meta_loss"]) - float(g["meta_loss0"])) < 2e-2
    assert abs(m.weights.eta.item() - float(g["eta0"])) < 1e-4
    assert abs(float(m.logged["avg_weight"]) - float(g["avg_weight0"])) < 2e-3
    grads = m.model.named_grads()
    names = [str(n) for n in g["names"]]
    assert sorted(names) == sorted(grads.keys())
    worst = 0.0
    for n in names:
        ref_norm = float(g["gnorm/" + n])
        got = grads[n]
        head = torch.tensor(g["ghead/" + n]).cuda()
        rel_norm = abs(got.norm().item() - ref_norm) / max(ref_norm, 1e-6)
        err = (got[:head.numel()] - head).abs().max().item()
        scale = max(head.abs().max().item(), ref_norm / got.numel() ** 0.5)
        worst = max(worst, rel_norm)
        assert rel_norm < 3e-2, f"{n}: |grad| {got.norm().item()} vs {ref_norm}"
        assert err < 1e-1 * scale + 1e-5, f"{n}: leading gradient entries off by {err} (scale {scale})"
    out1 = m.training_step(batch, 1)
    # the second step runs on the AdamW-updated weights and eta
    assert abs(float(out1["main_loss"]) - float(g["main_loss1"])) < 5e-2
    assert abs(float(out1["meta_loss"]) - float(g["meta_loss1"])) < 5e-2
    assert abs(m.weights.eta.item() - float(g["eta1"])) < 2e-4


@pytest.mark.parametrize("elide", [True, False])
def test_micro_batched_step_matches_golden_and_full_batch(golden_dir, elide):
    """hparams['micro_queries'] walks the batch in query groups (how T5-large 1+31 x 512 fits, BASELINE configs[3]):
    same losses / eta as the reference's golden step, gradients equal to the single-batch step's up to the fp32
    summation order of the split (bf16 rounding of each GEMM is per document, so it does not depend on the split)."""
    from pacednegatives_b200.lceT5 import LCEModel
    from pacednegatives_b200.scorer import T5Dims
    g, dims, hp, batch = load_case(golden_dir)
    w = ot.init_weights(dims, seed=int(g["seed"]))
    sd = {k: torch.as_tensor(v) for k, v in w.items()}
    Q = batch[0]["input_ids"].shape[0]
    assert Q >= 2
    full = LCEModel(dict(hp), dims=T5Dims(**dims.__dict__), state_dict=sd, elide_no_grad_forward=elide)
    micro = LCEModel(dict(hp, micro_queries=max(1, Q // 2)), dims=T5Dims(**dims.__dict__),
                     state_dict=sd, elide_no_grad_forward=elide)
    of, om = full.training_step(batch, 0), micro.training_step(batch, 0)
    assert abs(float(om["main_loss"]) - float(g["main_loss0"])) < 2e-2
    assert abs(float(om["main_loss"]) - float(of["main_loss"])) < 1e-5
    assert abs(float(om["meta_loss"]) - float(of["meta_loss"])) < 1e-5
    assert abs(micro.weights.eta.item() - full.weights.eta.item()) < 1e-7
    gf, gm = full.model.grads, micro.model.grads
    assert (gf - gm).norm().item() < 2e-3 * gf.norm().item()
    o1 = micro.training_step(batch, 1)
    assert abs(float(o1["main_loss"]) - float(g["main_loss1"])) < 5e-2
    assert abs(micro.weights.eta.item() - float(g["eta1"])) < 2e-4


def test_adamw_kernel_vs_transformers_rule():
    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    lib = _lib.lib()
    torch.manual_seed(3)
    n = 100003
    p = torch.randn(n, device="cuda")
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    rp, rm, rv = p.double().clone(), m.double().clone(), v.double().clone()
    lr, b1, b2, eps, wd = 1e-3, 0.9, 0.999, 1e-6, 0.01
    for step in range(1, 4):
        gr = torch.randn(n, device="cuda") * 0.1
        _lib.check(lib.pn_adamw(_lib.ptr(p), _lib.ptr(gr), _lib.ptr(m), _lib.ptr(v), n, lr, b1, b2, eps, wd, step,
                                0.5, _lib.stream()))
        gd = gr.double() * 0.5
        rm = b1 * rm + (1 - b1) * gd
        rv = b2 * rv + (1 - b2) * gd * gd
        rp = rp - lr * (1 - b2 ** step) ** 0.5 / (1 - b1 ** step) * rm / (rv.sqrt() + eps)
        rp = rp - lr * wd * rp
    assert (p.double() - rp).abs().max().item() < 1e-5


# ---------------------------------------------------------------------------------------------------------------
# The composed backward at the widths the bench runs.  Goldens: the reference's unmodified
# LCEModel.training_step on HF fp32 CPU (tests/golden/make_golden.py::gold_train_widths) at t5-base width
# (d 768, 12 heads, ff 3072), T5-large width (d 1024, 16 heads, ff 4096) — 2+2 layers, S 512, vocab 32128 — and
# SURVEY 8d cfg1 (the whole t5-small, Q 8, k 7, S 128, eta 0.3466).  Per gradient tensor: norm, relative L2 error
# over a seeded element sample (every element for tensors <= 65536), relative error of the row / column sums.
#
# Tolerances (bf16 operands, fp32 accumulation, against fp32 CPU; measured worst cases are written to
# gpurun_out/train_width_parity_*.json and copied to profiles/):
#   losses 2e-2 absolute x max(1, |loss| / 10): north_star's 2e-2 is stated for O(1) scores / losses; a random-init
#     32128-way CE is ~ 20, where one bf16 ulp of a logit is already 6e-2;
#   gradient tensors: relative L2 error <= 6e-2 for weight matrices (every entry is a sum over docs*S products of
#     two bf16-rounded factors: relative rounding 2^-9 each, summed incoherently, plus the bf16 dS / dP tiles of the
#     attention backward), <= 8e-2 for the RMSNorm gains and the relative-position tables (differences of large
#     cancelling sums); norms within 3e-2; row / column sums of each matrix within 2x the class tolerance of |grad|_F
#     (noise-scaled, see the loop).  Measured on the B200 (profiles/r4_train_width_parity.md): worst relative L2
#     0.036 (base), 0.040 (large), 0.047 (t5-small cfg1), norms within 0.8 %.
def _sample_index(name, numel, count=8192):
    import zlib
    if numel <= 65536:
        return np.arange(numel, dtype=np.int64)
    rng = np.random.RandomState(zlib.crc32(name.encode()) & 0x7FFFFFFF)
    return np.sort(rng.choice(numel, size=count, replace=False)).astype(np.int64)


def _load_wide(golden_dir, case):
    g = np.load(os.path.join(golden_dir, f"train_golden_{case}.npz"))
    v, d, kv, ff, L, Ld, H = [int(x) for x in g["dims"]]
    dims = ot.T5Dims(vocab=v, d_model=d, d_kv=kv, d_ff=ff, n_layers=L, n_dec_layers=Ld, n_heads=H)
    n, use_mean, lr, meta_lr, warm, total, eta = [float(x) for x in g["hp"]]
    hp = dict(n=int(n), use_mean=bool(use_mean), ignore_index=-100, lr=lr, meta_lr=meta_lr, warmup_steps=int(warm),
              total_steps=int(total), eta=eta)
    batch = ({"input_ids": torch.tensor(g["pids"]), "labels": torch.tensor(g["plab"])},
             {"input_ids": torch.tensor(g["nids"]), "labels": torch.tensor(g["nlab"])})
    return g, dims, hp, batch


def _tensor_class(name):
    if "layer_norm" in name:
        return "norm_gain"
    if "relative_attention_bias" in name:
        return "rel_bias"
    if name == "shared.weight":
        return "embedding"
    return "matrix"


WIDE_TOL = {"matrix": 6e-2, "embedding": 6e-2, "norm_gain": 8e-2, "rel_bias": 8e-2}


@pytest.mark.parametrize("case,xattn,cta_group", [
    ("base", True, 2), ("base", False, 2), ("base", True, 1),
    ("large", True, 2), ("large", False, 1),
    ("cfg1", True, 2), ("cfg1", False, 2)])
def test_training_step_at_bench_widths(golden_dir, case, xattn, cta_group):
    import json
    from pacednegatives_b200.lceT5 import LCEModel
    from pacednegatives_b200.scorer import T5Dims
    g, dims, hp, batch = _load_wide(golden_dir, case)
    w = ot.init_weights(dims, seed=int(g["seed"]))
    m = LCEModel(hp, dims=T5Dims(**dims.__dict__), state_dict={k: torch.as_tensor(v) for k, v in w.items()},
                 xattn=xattn, cta_group=cta_group)
    del w
    out = m.training_step(batch, 0)
    ref_main, ref_meta = float(g["main_loss0"]), float(g["meta_loss0"])
    ltol = 2e-2 * max(1.0, abs(ref_main) / 10.0)
    report = {"case": case, "xattn": xattn, "cta_group": cta_group,
              "main_loss": [float(out["main_loss"].detach()), ref_main],
              "meta_loss": [float(out["meta_loss"]), ref_meta],
              "eta": [m.weights.eta.item(), float(g["eta0"])], "tensors": {}}
    grads = m.model.named_grads()
    names = [str(n) for n in g["names"]]
    assert sorted(names) == sorted(grads.keys())
    worst = {}
    fails = []
    for n in names:
        got = grads[n]
        ref_norm = float(g["gnorm/" + n])
        idx = torch.from_numpy(_sample_index(n, got.numel())).cuda()
        ref = torch.tensor(g["gsample/" + n]).cuda().double()
        diff = (got[idx].double() - ref).norm().item()
        rel = diff / max(ref.norm().item(), 1e-30)
        rel_norm = abs(got.double().norm().item() - ref_norm) / max(ref_norm, 1e-30)
        entry = {"rel_l2": rel, "rel_norm": rel_norm}
        if "growsum/" + n in g.files:
            shape = (len(g["growsum/" + n]), len(g["gcolsum/" + n]))
            G = got.view(shape).double()
            for key, dim in (("growsum/", 1), ("gcolsum/", 0)):
                r = torch.tensor(g[key + n]).cuda().double()
                # Noise-scaled: independent entry errors of relative size eps move a vector of sums by
                # ~ eps * |G|_F whatever the sums themselves are (those of o.weight / wo.weight cancel to a small
                # fraction of |G|_F: the residual-stream gradient is nearly orthogonal to the all-ones vector), while a
                # coherent error (a head, a block, a diagonal of the bias gradient) moves them by up to sqrt(dim) more.
                entry[key[1:-1]] = ((G.sum(dim) - r).norm() / max(ref_norm, 1e-30)).item()
                entry[key[1:-1] + "_rel_to_sums"] = ((G.sum(dim) - r).norm() / max(r.norm().item(), 1e-30)).item()
        report["tensors"][n] = entry
        cls = _tensor_class(n)
        worst[cls] = max(worst.get(cls, 0.0), rel)
        if ref_norm < 1e-12:      # e.g. every loss above eta and nothing flows: both must be ~0
            if got.norm().item() > 1e-6:
                fails.append(f"{n}: reference gradient is zero, got norm {got.norm().item()}")
            continue
        if rel > WIDE_TOL[cls]:
            fails.append(f"{n}: relative L2 error {rel:.4f} > {WIDE_TOL[cls]}")
        if rel_norm > 3e-2:
            fails.append(f"{n}: |grad| off by {rel_norm:.4f}")
        for key in ("rowsum", "colsum"):
            if key in entry and entry[key] > 2 * WIDE_TOL[cls]:
                fails.append(f"{n}: {key} off by {entry[key]:.4f} of |grad|")
    report["worst_rel_l2_by_class"] = worst
    os.makedirs(os.path.join(os.path.dirname(golden_dir), "..", "gpurun_out"), exist_ok=True)
    with open(os.path.join(os.path.dirname(golden_dir), "..", "gpurun_out",
                           f"train_width_parity_{case}_x{int(xattn)}_c{cta_group}.json"), "w") as f:
        json.dump(report, f, indent=1)
    assert abs(report["main_loss"][0] - ref_main) < ltol, report["main_loss"]
    assert abs(report["meta_loss"][0] - ref_meta) < ltol, report["meta_loss"]
    assert abs(report["eta"][0] - report["eta"][1]) < 2e-4, report["eta"]
    assert abs(float(m.logged["avg_weight"]) - float(g["avg_weight0"])) < 5e-3
    assert not fails, "\n".join(fails[:20]) + f"\nworst by class: {worst}"
    if int(g["steps"]) > 1:   # the second step runs on the AdamW-updated weights
        out1 = m.training_step(batch, 1)
        # Adam's first update is lr * sign(g) per element, so the second step's loss also depends on the SIGN of
        # every small gradient entry: 3 % of the loss
        assert abs(float(out1["main_loss"]) - float(g["main_loss1"])) < 3e-2 * abs(float(g["main_loss1"]))
        assert abs(m.weights.eta.item() - float(g["eta1"])) < 5e-4

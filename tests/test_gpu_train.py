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
#     cancelling sums); norms within 3e-2; the Frobenius norms of every 64 x 64 block of each matrix (per-head
#     slices; no cancellation) within 2e-2 as a vector and 1.2e-1 for the worst block (relative to
#     its norm + 2 % of the largest block's: the decoder's q / k blocks at T = 2 are tiny and noisy).  Measured on the B200 (profiles/r4_train_width_parity.md): worst relative L2
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
    if "dropout" in g.files and float(g["dropout"][0]) > 0.0:
        hp["dropout_rate"], hp["seed"] = float(g["dropout"][0]), int(g["dropout"][1])
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
    _check_step_against_golden(golden_dir, case, xattn, cta_group, WIDE_TOL)


# Train-mode dropout (a10): the reference's unmodified training_step on HF in train() mode with dropout_rate 0.1, the
# masks of every dropout call replayed from the counter hash of csrc/pn_dropout.cuh (oracle/dropout.py), so the
# comparison is as exact as without dropout: both forwards of the step (different masks), the backward through the
# regenerated masks, eta, and the second step (seeds advance).  'ragged': S 77 / 90 (odd key pitch, positives and
# negatives as two model calls with document offsets).  Toy widths (d 128) get the toy-width bound 8e-2 (see the
# restricted-mode test below), the base width the bench-width bounds.
TOY_TOL = {"matrix": 8e-2, "embedding": 8e-2, "norm_gain": 1e-1, "rel_bias": 1e-1}


@pytest.mark.parametrize("case,cta_group", [("dropout_tiny", 2), ("dropout_tiny", 1), ("dropout_ragged", 2),
                                            ("dropout_base", 2)])
def test_training_step_with_dropout_vs_reference_golden(golden_dir, case, cta_group):
    _check_step_against_golden(golden_dir, case, True, cta_group, WIDE_TOL if case.endswith("base") else TOY_TOL)


def test_dropout_rate_zero_is_bit_identical_and_seeds_matter(golden_dir):
    """rate 0 goes through the same kernels with the masks off: identical bits to a model built without the option;
    at rate 0.1 two seeds give different losses, the same seed the same loss."""
    from pacednegatives_b200.lceT5 import LCEModel
    from pacednegatives_b200.scorer import T5Dims
    g, dims, hp, batch = load_case(golden_dir)
    sd = {k: torch.as_tensor(v) for k, v in ot.init_weights(dims, seed=int(g["seed"])).items()}

    def run(**extra):
        m = LCEModel(dict(hp, **extra), dims=T5Dims(**dims.__dict__), state_dict=sd)
        out = m.training_step(batch, 0)
        return float(out["main_loss"].detach()), m.model.grads.clone()
    l0, g0 = run()
    l1, g1 = run(dropout_rate=0.0)
    assert l0 == l1
    # weight gradients are accumulated with fp32 atomics (split-K): equal up to summation order
    assert (g0 - g1).abs().max().item() <= 1e-5 * g0.abs().max().item()
    la, _ = run(dropout_rate=0.1, seed=1)
    lb, _ = run(dropout_rate=0.1, seed=1)
    lc, _ = run(dropout_rate=0.1, seed=2)
    assert la == lb and abs(la - lc) > 1e-3 and abs(la - l0) > 1e-3


def _check_step_against_golden(golden_dir, case, xattn, cta_group, tol_by_class):
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
        if "gblock/" + n in g.files:
            ref_b = torch.tensor(g["gblock/" + n]).cuda().double()
            R, C = len(g["growsum/" + n]), len(g["gcolsum/" + n])
            G = got.view(R, C).double()
            gb = G[:R - R % 64, :C - C % 64].reshape(R // 64, 64, C // 64, 64).pow(2).sum((1, 3)).sqrt()
            # norms of the 64 x 64 blocks (per-head slices): free of cancellation, so independent entry errors of
            # relative size eps move them by ~ eps^2 / 2, a wrong head / block by O(1)
            entry["block_norms"] = ((gb - ref_b).norm() / max(ref_b.norm().item(), 1e-30)).item()
            entry["block_norm_worst"] = ((gb - ref_b).abs() / (ref_b + 2e-2 * ref_b.max())).max().item()
            # informational (both cancel heavily: sums over tokens of random-sign gradient rows)
            for key, dim in (("growsum/", 1), ("gcolsum/", 0)):
                r = torch.tensor(g[key + n]).cuda().double()
                entry[key[1:-1] + "_rel_to_sums"] = ((G.sum(dim) - r).norm() / max(r.norm().item(), 1e-30)).item()
        report["tensors"][n] = entry
        cls = _tensor_class(n)
        worst[cls] = max(worst.get(cls, 0.0), rel)
        if ref_norm < 1e-12:      # e.g. every loss above eta and nothing flows: both must be ~0
            if got.norm().item() > 1e-6:
                fails.append(f"{n}: reference gradient is zero, got norm {got.norm().item()}")
            continue
        if rel > tol_by_class[cls]:
            fails.append(f"{n}: relative L2 error {rel:.4f} > {tol_by_class[cls]}")
        if rel_norm > 3e-2:
            fails.append(f"{n}: |grad| off by {rel_norm:.4f}")
        if "block_norms" in entry and entry["block_norms"] > 2e-2:
            fails.append(f"{n}: 64x64 block norms off by {entry['block_norms']:.4f}")
        if "block_norm_worst" in entry and entry["block_norm_worst"] > 1.2e-1:
            fails.append(f"{n}: one 64x64 block's norm off by {entry['block_norm_worst']:.4f}")
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


# ---------------------------------------------------------------------------------------------------------------
# mode='restricted' (north_star items (1) + (2)): two-row head + group log-softmax LCE, backward from dz2
@pytest.mark.parametrize("xattn", [True, False])
def test_restricted_mode_step_vs_oracle_and_hf_autograd(golden_dir, xattn):
    """Loss against oracle/loss.py::lce_restricted on the oracle's fp32 z2; every gradient tensor against torch
    autograd through HF T5 (fp32, CPU) of the same objective: mean_q [logsumexp_j s_qj - s_q0],
    s = log_softmax(logits[:, 0, (rel, nrel)])[:, 0]."""
    from oracle import loss as ol
    from pacednegatives_b200.lceT5 import LCEModel
    from pacednegatives_b200.scorer import T5Dims
    from transformers import T5Config, T5ForConditionalGeneration
    g, dims, hp, batch = load_case(golden_dir)
    rel, nrel = 117, 213
    w = ot.init_weights(dims, seed=int(g["seed"]))
    hp = dict(hp, mode="restricted", eta=1.0)
    m = LCEModel(hp, dims=T5Dims(**dims.__dict__), state_dict={k: torch.as_tensor(v) for k, v in w.items()},
                 xattn=xattn, rel_id=rel, nrel_id=nrel)
    out = m.training_step(batch, 0)
    pids, nids = g["pids"], g["nids"]
    Q, k = pids.shape[0], hp["n"]
    _, zp = ot.score(w, dims, pids, rel, nrel)
    _, zn = ot.score(w, dims, nids, rel, nrel)
    z = np.concatenate([zp[:, None], zn.reshape(Q, k, 2)], axis=1)
    eloss, _ = ol.lce_restricted(z.reshape(-1, 2), 1 + k)
    assert abs(float(out["main_loss"].detach()) - float(eloss.mean())) < 2e-2
    # HF autograd of the same objective
    cfg = T5Config(vocab_size=dims.vocab, d_model=dims.d_model, d_kv=dims.d_kv, d_ff=dims.d_ff,
                   num_layers=dims.n_layers, num_decoder_layers=dims.n_dec_layers, num_heads=dims.n_heads,
                   dropout_rate=0.0, decoder_start_token_id=0, pad_token_id=0, eos_token_id=1,
                   feed_forward_proj="relu", tie_word_embeddings=True)
    hf = T5ForConditionalGeneration(cfg).eval()
    sd = {k_: torch.tensor(v) for k_, v in w.items()}
    sd["encoder.embed_tokens.weight"] = sd["decoder.embed_tokens.weight"] = sd["lm_head.weight"] = sd["shared.weight"]
    hf.load_state_dict(sd, strict=False)
    ids = torch.tensor(np.concatenate([pids, nids]))
    lg = hf(input_ids=ids, decoder_input_ids=torch.zeros(len(ids), 1, dtype=torch.long)).logits[:, 0, [rel, nrel]]
    s = torch.log_softmax(lg, dim=1)[:, 0]
    sg = torch.cat([s[:Q, None], s[Q:].view(Q, k)], dim=1)
    ref_loss = (torch.logsumexp(sg, dim=1) - sg[:, 0]).mean()
    ref_loss.backward()
    assert abs(float(out["main_loss"].detach()) - float(ref_loss)) < 2e-2
    grads = m.model.named_grads()
    # Tolerances (relative L2 per tensor, measured values in the assertion message on failure):
    #   encoder tensors 8e-2 at this toy width (d 128: each entry sums few products, so bf16 rounding averages out
    #   less than at the bench widths, where the same kernels measure 0.036 - 0.047); measured 0.062 worst.
    #   decoder tensors 2e-1: with ONE decoder step every document feeds the decoder the same start token, so e.g.
    #   the gradient of decoder.block.0 v.weight is the rank-one (sum_r dY_r) h^T, and the LCE gradient sums to zero
    #   over each group (softmax minus one-hot): the per-document terms cancel to a fraction of their size and the
    #   bf16 rounding of each dY_r (2^-9 relative) is amplified by that factor; measured 0.095 worst.
    errs, fails = {}, []
    for n_, p_ in hf.named_parameters():
        ref = p_.grad.reshape(-1).cuda()
        got = grads[n_]
        rn = ref.norm().item()
        if rn < 1e-10:
            if got.norm().item() > 1e-6:
                fails.append(f"{n_}: reference gradient is zero, got {got.norm().item():.3e}")
            continue
        errs[n_] = (got - ref).norm().item() / rn
        tol = 2e-1 if n_.startswith("decoder.") else 8e-2
        if errs[n_] > tol:
            fails.append(f"{n_}: relative L2 error {errs[n_]:.4f} > {tol} (|grad| {rn:.3e})")
    worst = sorted(errs.items(), key=lambda kv: -kv[1])[:6]
    assert not fails, "\n".join(fails) + f"\nworst: {worst}"
    # the bulk of the gradient (all tensors as one vector) is far tighter than the worst small tensor
    tot_ref = torch.cat([p_.grad.reshape(-1) for _, p_ in hf.named_parameters()]).cuda()
    tot_got = torch.cat([grads[n_] for n_, _ in hf.named_parameters()])
    assert (tot_got - tot_ref).norm().item() / tot_ref.norm().item() < 5e-2, worst


def test_fused_norm_backward_matches_the_separate_pass(golden_dir):
    """"fuse_norm_bwd" (default; RMSNorm backward inside the encoder's data-gradient GEMMs, row dot taken from the
    kernels that produce the output gradient) against the separate norm_bwd pass ("fuse_norm_bwd" = 0) on the same
    step at t5-small width (cfg1 golden inputs): same loss bits (the forward is shared), every gradient tensor within
    1e-2 relative L2 — the two differ only in where the row dot is rounded (bf16 G . h against fp32 Ys . y) and in
    dh taking G unrounded instead of through bf16."""
    from pacednegatives_b200.lceT5 import LCEModel
    from pacednegatives_b200.scorer import T5Dims
    g, dims, hp, batch = _load_wide(golden_dir, "cfg1")
    sd = {k: torch.as_tensor(v) for k, v in ot.init_weights(dims, seed=int(g["seed"])).items()}

    def run(fuse):
        m = LCEModel(hp, dims=T5Dims(**dims.__dict__), state_dict=sd, cta_group=2, options={"fuse_norm_bwd": fuse})
        out = m.training_step(batch, 0)
        return float(out["main_loss"].detach()), {k: v.clone() for k, v in m.model.named_grads().items()}
    l1, g1 = run(1)
    l0, g0 = run(0)
    assert l0 == l1
    worst = 0.0
    for n in g0:
        ref = g0[n].double()
        if ref.norm().item() < 1e-12:
            continue
        worst = max(worst, ((g1[n].double() - ref).norm() / ref.norm()).item())
    assert worst < 1e-2, worst

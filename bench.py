#!/usr/bin/env python
"""bench.py — the reference's headline metrics on B200 (BASELINE.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload all|rerank|train|select] [--impl reference]

BASELINE.json's metric has two halves, and the default workload "all" runs both in one process:
  * the JSON line's metric / value / ms_per_step / e2e / roofline are the RE-RANK half (BASELINE configs[1]):
    monoT5-base re-rank of synthetic BM25 top-1000 lists, seq 512, bf16 tensor-core GEMMs with fp32 accumulation.
    One step = `--queries-per-step` queries x 1000 candidate prompts scored (encoder + one decoder step + two-row
    head), ranked and evaluated (top-10, nDCG@10, RR) — SURVEY 8 a1 + a9 + a12; at N > 1 the per-rank score tables
    are all-gathered inside the step.  metric = pairs/s over all ranks.
  * the TRAINING half (configs[2]: monoT5-base LCE step, 16 queries x (1+15) x 512 per rank, NCCL gradient
    all-reduce at N > 1) is reported as flat scalars `config.train_*` (and its whole line on stderr), plus a short
    pass over the selection kernels (configs[4]) as `config.select_*`.

workload "rerank" / "train" / "select": one half alone, as the line's metric (`--model large` = configs[3]).
`--workload select --sweep`: the SURVEY 8d cfg5 sweep (M x eta x levels x layouts) as a table on the line.

Multi-GPU: launched by torchrun, one rank per GPU, queries sharded by rank; weak scaling.
`--impl reference` times the reference's own CPU code path for the same metric (HuggingFace T5 fp32
forward + the monoT5 score formula + ranking/metrics on the host cores).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


RANK, LOCAL_RANK, WORLD = env_int("RANK", 0), env_int("LOCAL_RANK", 0), env_int("WORLD_SIZE", 1)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "bf16_burst": d.get("bf16_tflops", 1590.0),
                "bf16_sustained": d.get("bf16_tflops_sustained", 1400.0), "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "src": "fallback"}


def ncu_traffic(workload, rows_per_launch):
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture
    (profiles/traffic.json), scaled by rows when the bench launch processes a different number of token rows than
    the captured one (every operand and output of these GEMMs but the weights scales with rows).  None when no
    capture of that workload is committed."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None, None
    with open(p) as f:
        d = json.load(f).get(workload)
    if not d:
        return None, None
    return d["traffic_bytes_per_launch"] * rows_per_launch / d["rows_per_launch"], d["source"]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._pump, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)),
                "power_w_max": float(max(pw)), "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# synthetic workloads
def synth_ids(n_docs, S, seed, vocab_hi=32100):
    """Prompt token ids ~ U{2..32099}, last column = 1 (</s>) — cfg2 of SURVEY §8d."""
    rng = np.random.RandomState(seed)
    ids = rng.randint(2, vocab_hi, size=(n_docs, S)).astype(np.int32)
    ids[:, -1] = 1
    return ids


def synth_rels(Q, C, seed):
    rng = np.random.RandomState(seed)
    return rng.choice(4, size=(Q, C), p=[0.95, 0.03, 0.015, 0.005]).astype(np.uint8)


# stdout carries exactly one JSON line.  Native libraries write there too (NCCL prints "NCCL version ..." on stdout
# at the box's debug level), so file descriptor 1 is pointed at stderr for the whole run and the line goes to the
# saved descriptor.
_STDOUT_FD = None


def claim_stdout():
    global _STDOUT_FD
    if _STDOUT_FD is None:
        sys.stdout.flush()
        _STDOUT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    text = (json.dumps(line) + "\n").encode()
    if _STDOUT_FD is None:
        sys.stdout.write(text.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_STDOUT_FD, text)


def dist_setup(use_cuda):
    claim_stdout()
    if WORLD > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl" if use_cuda else "gloo", rank=RANK, world_size=WORLD,
                                device_id=torch.device("cuda", LOCAL_RANK) if use_cuda else None)
        return dist
    return None


def barrier(dist):
    if dist is not None:
        dist.barrier()


def max_over_ranks(dist, ms, device):
    if dist is None:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed_region(dist, fn, steps):
    """barrier + synchronize on both sides, CUDA events on the launching (current) stream."""
    torch.cuda.synchronize()
    barrier(dist)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    barrier(dist)
    return max_over_ranks(dist, e0.elapsed_time(e1), "cuda")


def hbm_gemm_entry(h_ms, h_n, h_bytes, step_ms_total, peaks):
    """GEMM launches under ~200 FLOP per compulsory byte (skinny cross-attention products, the fp32 residual
    read-modify-write of the narrow output projections): bounded by HBM, reported against the HBM peak."""
    gbs = h_bytes / (h_ms / 1e3) / 1e9 if h_ms else 0.0
    return {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
            "frac": gbs / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None, "launches": int(h_n),
            "share_of_step": h_ms / step_ms_total if step_ms_total else None}


def prof_read(kind):
    import ctypes as C
    from pacednegatives_b200 import _lib
    ms, n, work = C.c_double(), C.c_int64(), C.c_double()
    _lib.check(_lib.lib().pn_prof_read(kind, C.byref(ms), C.byref(n), C.byref(work)))
    return ms.value, n.value, work.value


# ------------------------------------------------------------------------------------------------
# CPU / library baselines
def _hf_t5(large=False, train=False, dropout=0.0):
    from transformers import T5Config, T5ForConditionalGeneration
    d, ff, L, H = (1024, 4096, 24, 16) if large else (768, 3072, 12, 12)
    cfg = T5Config(vocab_size=32128, d_model=d, d_kv=64, d_ff=ff, num_layers=L, num_decoder_layers=L,
                   num_heads=H, dropout_rate=dropout, decoder_start_token_id=0, pad_token_id=0, eos_token_id=1)
    torch.manual_seed(0)
    m = T5ForConditionalGeneration(cfg)
    return m.train() if train else m.eval()


def cpu_rerank_fn(n_docs, S):
    """The reference's CPU path for re-ranking n_docs prompts: HuggingFace T5ForConditionalGeneration fp32 forward
    (what lceT5.py:132 / pyterrier_t5 call) when transformers is importable, else the oracle's torch-CPU
    restatement; then the monoT5 score formula (distill/wrapper.py:34-38).  Returns (run(ids) -> scores, how)."""
    try:
        model = _hf_t5()
        dec = torch.zeros(n_docs, 1, dtype=torch.long)

        def run(x):
            with torch.no_grad():
                lg = model(input_ids=x, decoder_input_ids=dec[: x.shape[0]]).logits
                return torch.log_softmax(lg[:, 0, [1176, 6136]], dim=1)[:, 0]
        return run, "transformers T5ForConditionalGeneration fp32 eager CPU + monoT5 score (distill/wrapper.py:34-38)"
    except Exception:  # transformers missing or incompatible: the oracle's restatement
        from oracle import t5 as ot
        w = ot.init_weights(ot.T5_BASE, seed=0)

        def run(x):
            return torch.as_tensor(ot.score(w, ot.T5_BASE, x.numpy(), 1176, 6136)[0])
        return run, "oracle.t5 torch-CPU fp32 restatement"


def cpu_rerank_sample(n_docs, S, threads=None):
    if threads:
        torch.set_num_threads(threads)
    ids = torch.from_numpy(synth_ids(n_docs, S, 999).astype(np.int64))
    run, how = cpu_rerank_fn(8, S)
    bs = 8
    run(ids[:bs])  # warm-up
    t0 = time.perf_counter()
    for lo in range(0, n_docs, bs):
        run(ids[lo:lo + bs])
    dt = time.perf_counter() - t0
    return n_docs / dt, how, dt


def hf_train_sample(n_docs, S, steps, warmup, threads=None, large=False, device="cpu", dtype=torch.float32):
    """The reference's path for one training step's model work on a bounded sample: HF
    T5ForConditionalGeneration forward + backward + AdamW over n_docs prompts with 2 label tokens and the
    reference's loss (lceT5.py:149-156).  Returns (docs/s, description, seconds)."""
    if threads:
        torch.set_num_threads(threads)
    model = _hf_t5(large, train=True).to(device=device, dtype=dtype)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3, eps=1e-6, weight_decay=0.0)
    ids = torch.from_numpy(synth_ids(n_docs, S, 4567).astype(np.int64)).to(device)
    labels = torch.tensor([[1176, 1]] + [[6136, 1]] * (n_docs - 1)).to(device)
    loss_fn = torch.nn.CrossEntropyLoss(ignore_index=-100, reduction="none")

    def run():
        logits = model(input_ids=ids, labels=labels).logits
        ce = loss_fn(logits.view(-1, logits.size(-1)).float(), labels.view(-1))
        opt.zero_grad()
        ce.mean().backward()
        opt.step()
        return ce
    for _ in range(warmup):
        run()
    if device != "cpu":
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        run()
    if device != "cpu":
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    how = f"transformers T5ForConditionalGeneration {str(dtype).split('.')[-1]} eager {device} fwd+bwd+AdamW"
    return n_docs * steps / dt, how, dt


def library_baseline(S=512, k=15):
    """SURVEY 2 / BASELINE.md 4 row 6: what the reference would actually run on this GPU — the same HF eager
    path (cuBLAS / ATen kernels, none of this repo's) in bf16 on the same B200.  Re-rank: 64 docs x 512 forward +
    score; train: fwd + bwd + AdamW on 32 docs x 512 with 2 label tokens.  Flat scalars for `config`."""
    out = {}
    try:
        model = _hf_t5().to(device="cuda", dtype=torch.bfloat16)
        ids = torch.from_numpy(synth_ids(64, S, 999).astype(np.int64)).cuda()
        dec = torch.zeros(64, 1, dtype=torch.long, device="cuda")

        def run():
            with torch.no_grad():
                lg = model(input_ids=ids, decoder_input_ids=dec).logits
                return torch.log_softmax(lg[:, 0, [1176, 6136]].float(), dim=1)[:, 0]
        for _ in range(3):
            run()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            run()
        e1.record()
        torch.cuda.synchronize()
        out["library_baseline_rerank_pairs_per_s"] = 64 * 5 / (e0.elapsed_time(e1) / 1e3)
        out["library_baseline_rerank"] = "HF T5ForConditionalGeneration bf16 eager on this GPU, 64 docs x 512 per call"
        del model, ids
        torch.cuda.empty_cache()
        v, how, dt = hf_train_sample(32, S, 3, 2, device="cuda", dtype=torch.bfloat16)
        out["library_baseline_train_queries_per_s"] = v / (1 + k)
        out["library_baseline_train"] = f"{how}, 32 docs x 512 per step; docs/s / {1 + k}"
        torch.cuda.empty_cache()
    except Exception as e:   # say so rather than invent a number
        out["library_baseline_error"] = f"{type(e).__name__}: {e}"[:200]
    return out


# ------------------------------------------------------------------------------------------------
# workload: re-rank (BASELINE.json configs[1])
def run_rerank(args, dist):
    from pacednegatives_b200 import _lib
    from pacednegatives_b200.scorer import MonoT5Scorer, T5_BASE, random_state_dict, flops_per_doc
    lib = _lib.lib()
    S, C, K = 512, 1000, 10
    Qs = args.queries_per_step
    n_docs = Qs * C
    dims = T5_BASE
    scorer = MonoT5Scorer(dims, random_state_dict(dims, seed=0), chunk_docs=args.chunk_docs,
                          cta_group=args.cta_group)
    total = args.steps + args.warmup
    nbuf = min(total, 4)   # distinct step inputs cycled; each is 16 MB of ids feeding > 1 GB of activations
    host_ids = [torch.from_numpy(synth_ids(n_docs, S, 2345 + 1000 * RANK + i)).pin_memory()
                for i in range(nbuf)]
    dev_ids = [h.cuda() for h in host_ids]
    rels_h = torch.from_numpy(synth_rels(Qs, C, 3456 + RANK)).pin_memory()
    docids_h = torch.arange(Qs * C, dtype=torch.int32).view(Qs, C).pin_memory()
    rels_d, docids_d = rels_h.cuda(), docids_h.cuda()
    out_host = {"top_ids": torch.empty(Qs, K, dtype=torch.int32).pin_memory(),
                "top_scores": torch.empty(Qs, K, dtype=torch.float32).pin_memory(),
                "metrics": torch.empty(Qs, 4, dtype=torch.float32).pin_memory()}
    mean_host = torch.empty(4, dtype=torch.float32).pin_memory()
    # N > 1: the per-rank [Q, C] score tables (+ top-10, measures) are all-gathered and the global means
    # all-reduced INSIDE the step (SURVEY 8e collective 3; eval.py:23-26 evaluates one table)
    gather = "scores" if dist is not None else "none"

    def step_dev(i):
        scorer.rerank(dev_ids[i % nbuf].view(Qs, C, S), docids_d, rels_d, K, gather=gather)

    def step_e2e(i):
        ids = host_ids[i % nbuf].cuda(non_blocking=True)
        out = scorer.rerank(ids.view(Qs, C, S), docids_h.cuda(non_blocking=True),
                            rels_h.cuda(non_blocking=True), K, gather=gather)
        for k, v in out_host.items():
            v.copy_(out[k], non_blocking=True)
        if "mean_metrics" in out:
            mean_host.copy_(out["mean_metrics"], non_blocking=True)
        torch.cuda.current_stream().synchronize()   # the caller reads the step's ranking and metrics

    for i in range(args.warmup):
        step_dev(i)
    clocks = ClockSampler(LOCAL_RANK)
    clocks.start()
    lib.pn_prof_enable(0)
    l0 = _lib.launch_count()
    ms = timed_region(dist, step_dev, args.steps)          # `value`: no per-launch events in this region
    launches = _lib.launch_count() - l0
    clk = clocks.stop()
    # separate pass for the per-class breakdown (CUDA events around every launch of a class perturb the step)
    psteps = max(1, min(args.steps, 3))
    lib.pn_prof_enable(1)
    ms_prof = timed_region(dist, step_dev, psteps)
    lib.pn_prof_enable(0)
    g_ms, g_n, g_flops = prof_read(0)
    a_ms, a_n, a_flops = prof_read(1)
    h_ms, h_n, h_bytes = prof_read(5)
    for i in range(max(1, min(args.warmup, 2))):
        step_e2e(i)
    e2e_ms = timed_region(dist, step_e2e, args.steps)

    pairs = n_docs * args.steps * WORLD
    value = pairs / (ms / 1e3)
    e2e_value = pairs / (e2e_ms / 1e3)
    peaks = measured_peaks()
    fpd = flops_per_doc(dims, S, 1, False)
    traffic, traffic_src = ncu_traffic("rerank", min(args.chunk_docs, n_docs) * S)
    gemm_tflops = g_flops / (g_ms / 1e3) / 1e12 if g_ms > 0 else 0.0
    attn_tflops = a_flops / (a_ms / 1e3) / 1e12 if a_ms else 0.0
    hbm = hbm_gemm_entry(h_ms, h_n, h_bytes, ms_prof, peaks)
    gather_bytes = (Qs * C * 4 + Qs * K * 8 + Qs * 16) * WORLD if dist is not None else 0
    line = {
        "metric": "monoT5-base rerank pairs/s", "value": value, "unit": "pairs/s", "n_gpus": WORLD,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic (random-init t5-base dims, ids U{2..32099}, graded qrels)",
        "config": {"workload": "monoT5-base re-rank, BM25 top-1000, seq 512 (BASELINE configs[1])",
                   "queries_per_step_per_gpu": Qs, "candidates": C, "seq_len": S,
                   "chunk_docs": args.chunk_docs, "cta_group": args.cta_group,
                   "l2": "inputs+activations per step (>1 GB) exceed the 126 MB L2; no flush needed",
                   "parallelism": (f"queries sharded over {WORLD} rank(s); all_gather of the [Q,1000] score tables, "
                                   f"top-10 and measures + 5-float all-reduce of the means inside every step"
                                   if dist is not None else "1 rank, no collective"),
                   "score_gather_bytes_per_step": int(gather_bytes),
                   "timing": "value: per-launch profiling events OFF; class breakdown from a separate "
                             f"{psteps}-step pass ({ms_prof / psteps:.1f} ms/step with events)"},
        "gpu_launches": int(launches),
        "e2e": {"value": e2e_value, "unit": "pairs/s",
                "h2d_bytes_per_step": int(host_ids[0].numel() * 4 + rels_h.numel() + docids_h.numel() * 4),
                "d2h_bytes_per_step": int(sum(v.numel() * v.element_size() for v in out_host.values())
                                          + (16 if dist is not None else 0)),
                "ms_per_step": e2e_ms / args.steps},
        "roofline": {"bound": "tensor",
                     "kernel": "gemm_bf16_kernel (tcgen05; tensor-bound T5 linears, HBM-bound launches listed apart)",
                     "achieved": gemm_tflops, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
                     "frac": gemm_tflops / peaks["bf16_sustained"], "peak_source": peaks["src"] + " sustained",
                     "launches": int(g_n), "avg_launch_ms": g_ms / max(g_n, 1),
                     "share_of_step": g_ms / ms_prof if ms_prof else None,
                     "traffic": traffic, "traffic_source": traffic_src,
                     # flat copies of the other kernel classes and of the whole-model rate (SURVEY 8d count)
                     "model_tflops": value * fpd / 1e12 / WORLD,
                     "model_frac": value * fpd / 1e12 / WORLD / peaks["bf16_sustained"],
                     "attn_tflops": attn_tflops, "attn_launches": int(a_n),
                     "attn_share": a_ms / ms_prof if ms_prof else None,
                     "attn_frac": attn_tflops / peaks["bf16_sustained"],
                     "hbm_gemm_gbs": hbm["achieved"], "hbm_gemm_frac": hbm["frac"],
                     "hbm_gemm_share": hbm["share_of_step"], "hbm_gemm_launches": hbm["launches"]},
        "clocks": clk,
    }
    line["model_tflops"] = line["roofline"]["model_tflops"]
    line["model_frac_of_sustained_peak"] = line["roofline"]["model_frac"]
    del scorer, dev_ids
    torch.cuda.empty_cache()
    return line


# ------------------------------------------------------------------------------------------------
# workload: paced selection + weighting bandwidth kernels (BASELINE.json configs[4])
def _time_kernel(fn, steps, warmup, flush):
    for _ in range(max(warmup, 3)):
        fn()
    ts = []
    for _ in range(steps):
        flush.zero_()     # inputs that fit the 126 MB L2 would otherwise be read from cache
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sum(ts) / len(ts)


def cpu_select_baseline(budget_s=20.0):
    """SURVEY 8d CPU baseline of the selection workload: the reference-shaped Python loops (oracle/baseline.py)
    on one host core each — LCEWeights.forward at M in {64, 1024, 16384}, LCEDataset.sample x 1000,
    TripletDataset.get_items x 100000, and the trec_eval restatement in NumPy for 200 queries x 1000."""
    from oracle import baseline as ob, metrics as om
    rng = np.random.RandomState(5678)
    out = {}
    eta = torch.nn.Parameter(torch.tensor([0.3466]))
    per_elem = None
    for M in (64, 1024, 16384):
        loss = torch.from_numpy(rng.exponential(0.7, M).astype(np.float32))
        t0 = time.perf_counter()
        ob.lce_weights_loop(loss, eta)
        dt = time.perf_counter() - t0
        out[f"weights_loop_M{M}_s"] = dt
        per_elem = dt / M
    out["weights_loop_elems_per_s"] = 1.0 / per_elem
    C, k = 1000, 15
    neg_idx = [((q * 7919 + np.arange(C) * 104729) % 8841823).astype(np.int64) for q in range(64)]
    np.random.seed(0)
    t0 = time.perf_counter()
    ob.sample_loop(neg_idx, k, 0.3466, 0.01, 1000)
    out["sample_queries_per_s"] = 1000 / (time.perf_counter() - t0)
    wts = rng.rand(100000).tolist()
    t0 = time.perf_counter()
    ob.get_items_loop(neg_idx, wts)
    out["get_items_queries_per_s"] = 100000 / (time.perf_counter() - t0)
    Qm = 200
    sc = rng.randn(Qm, C).astype(np.float32)
    rl = (rng.rand(Qm, C) > 0.95).astype(np.uint8)
    di = np.stack(neg_idx)[np.arange(Qm) % 64].astype(np.int32)
    t0 = time.perf_counter()
    om.batch_metrics(sc, di, rl, 10)
    out["metrics_queries_per_s"] = Qm / (time.perf_counter() - t0)
    # one pass of the GPU bench's kernel list per query, as the reference would run it: weights (1 element per
    # query), one level pick, one binomial sample, metrics for the 20 % of queries the bench evaluates
    per_query = (1.0 / out["weights_loop_elems_per_s"] + 2.0 / out["get_items_queries_per_s"]
                 + 1.0 / out["sample_queries_per_s"] + 0.2 / out["metrics_queries_per_s"])
    out["queries_per_s"] = 1.0 / per_query
    return out


def run_select(args, dist, quick=False):
    from pacednegatives_b200 import _lib
    from pacednegatives_b200 import weights as W, loss as L, metrics as M
    from pacednegatives_b200.dataloader import CandidateMatrix, ROW
    dev = torch.device("cuda", LOCAL_RANK)
    Q, C, k = (args.select_queries if not quick else min(args.select_queries, 1 << 18)), 1000, 15
    steps = args.steps if not quick else min(args.steps, 5)
    peaks = measured_peaks()
    # cand[q, r] = (q*7919 + r*104729) mod 8841823, built on the device in slabs
    cand = torch.empty(Q, C, dtype=torch.int32, device=dev)
    r = torch.arange(C, device=dev, dtype=torch.int64) * 104729
    for lo in range(0, Q, 1 << 16):
        hi = min(Q, lo + (1 << 16))
        q = torch.arange(lo + RANK * Q, hi + RANK * Q, device=dev, dtype=torch.int64) * 7919
        cand[lo:hi] = ((q[:, None] + r[None, :]) % 8841823).to(torch.int32)
    cm_row = CandidateMatrix(cand, ROW)
    cm_rank = cm_row.rank_major()
    g = torch.Generator(device=dev).manual_seed(5678 + RANK)
    Mw = args.select_elems if not quick else min(args.select_elems, 1 << 26)
    losses = torch.empty(Mw, device=dev).exponential_(1 / 0.7, generator=g)
    eta = torch.tensor([0.3466], device=dev)
    wq = torch.rand(Q, device=dev, generator=g)
    table, _ = cm_row.binomial_table(0.3466, 0.01)
    G = 16
    z2 = torch.randn(Q * G, 2, device=dev, generator=g)
    Qm = min(Q, 200000)
    scores = torch.randn(Qm, C, device=dev, generator=g)
    rels = (torch.rand(Qm, C, device=dev, generator=g) > 0.95).to(torch.uint8)
    docids = cand[:Qm]
    nt = 1 << 24
    row_tab = torch.ones(nt, device=dev)
    ce_p = torch.rand(nt, device=dev, generator=g)
    ce_n = torch.rand(nt, device=dev, generator=g)
    lib = _lib.lib()

    def k_weights():
        W.spl_weights(losses, eta, 0.0, 0, want_keep=True, want_stats=True)

    def k_table():
        _lib.check(lib.pn_weights_table_update(_lib.ptr(row_tab), _lib.ptr(ce_p), _lib.ptr(ce_n), row_tab.numel(),
                                               1.0 / 16, 0.1, _lib.stream()))

    kernels = [
        ("spl_weights (eta threshold, keep mask, stats)", k_weights, Mw * 9.0),
        ("weight table update", k_table, row_tab.numel() * 16.0),
        ("level select, scalar level, rank-major", lambda: cm_rank.select_level(0.5, check=False), Q * 8.0),
        ("level select, per-query weight, row-major gather", lambda: cm_row.select_items(wq, check=False), Q * 16.0),
        ("binomial select k=15 (Philox stream)", lambda: cm_row.select_binomial(table, k, seed=7),
         Q * k * 12.0 + 4.0 * Q),
        ("restricted LCE fwd (G=16)", lambda: L.lce_restricted(z2, G), Q * (G * 8.0 + 4.0 + G * 4.0)),
        ("top-k + nDCG@10/RR (C=1000)", lambda: M.topk_metrics(scores, docids, rels, 10), Qm * (C * 9.0 + 10 * 8 + 16)),
    ]
    res = []
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    l0 = _lib.launch_count()
    tot_ms = 0.0
    for name, fn, nbytes in kernels:
        ms = _time_kernel(fn, steps, args.warmup, flush)
        tot_ms += ms
        gbs = nbytes / (ms / 1e3) / 1e9
        res.append({"kernel": name, "ms": ms, "algorithmic_bytes": nbytes, "achieved_gbs": gbs,
                    "frac_of_hbm_peak": gbs / peaks["hbm_gbs"]})
    launches = _lib.launch_count() - l0
    sweep = None
    if args.sweep and not quick:
        # M in {2^20, 2^24, 2^28} x eta in {0.1, 0.3466, 0.5, 1.0} (weighting); levels 1..10 in both
        # candidate layouts (scalar level and per-query weights)
        sweep = []
        for Msz in (1 << 20, 1 << 24, 1 << 28):
            if Msz > Mw:
                continue
            sub = losses[:Msz]
            for e in (0.1, 0.3466, 0.5, 1.0):
                et = torch.tensor([e], device=dev)
                ms = _time_kernel(lambda: W.spl_weights(sub, et, 0.0, 0, want_keep=True, want_stats=True),
                                  steps, 3, flush)
                sweep.append({"kernel": "spl_weights", "M": Msz, "eta": e, "ms": ms,
                              "gbs": Msz * 9.0 / ms / 1e6, "frac": Msz * 9.0 / ms / 1e6 / peaks["hbm_gbs"]})
        for lvl in range(1, 11):
            w_l = lvl / 10.0
            for lname, cm in (("rank-major", cm_rank), ("row-major", cm_row)):
                ms = _time_kernel(lambda: cm.select_level(w_l, check=False), steps, 3, flush)
                # row-major scalar level: one 32-byte sector per query is moved for 4 useful bytes
                sweep.append({"kernel": "level select (scalar level)", "layout": lname, "level": lvl, "ms": ms,
                              "gbs": Q * 8.0 / ms / 1e6, "frac": Q * 8.0 / ms / 1e6 / peaks["hbm_gbs"]})
            wl = (wq * 0 + w_l).contiguous()
            ms = _time_kernel(lambda: cm_row.select_items(wl, check=False), steps, 3, flush)
            sweep.append({"kernel": "level select (per-query weight)", "layout": "row-major", "level": lvl, "ms": ms,
                          "gbs": Q * 16.0 / ms / 1e6, "frac": Q * 16.0 / ms / 1e6 / peaks["hbm_gbs"]})
    top = res[0]
    sel_traffic, sel_traffic_src = ncu_traffic("select", Mw)   # dram bytes of one spl_weights launch, scaled by elements
    line = {"metric": "paced selection + weighting kernels, queries/s through one pass of all kernels",
            "value": Q * WORLD / (tot_ms / 1e3), "unit": "queries/s", "n_gpus": WORLD, "steps": steps,
            "warmup": max(args.warmup, 3), "ms_per_step": tot_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32/i32/f64", "data": "synthetic (SURVEY cfg5 generators)",
            "config": {"workload": "paced negative sampling + weighting kernel sweep (BASELINE configs[4])",
                       "queries": Q, "candidates": C, "k": k, "weight_elems": Mw, "l2": "256 MB flush between iterations",
                       "parallelism": f"{WORLD} rank(s), queries sharded, no collective"},
            "gpu_launches": int(launches),
            "e2e": None,
            "roofline": {"bound": "hbm", "kernel": top["kernel"], "achieved": top["achieved_gbs"], "peak": peaks["hbm_gbs"],
                         "unit": "GB/s", "frac": top["frac_of_hbm_peak"], "peak_source": peaks["src"],
                         "traffic": sel_traffic, "traffic_source": sel_traffic_src},
            "kernels": res}
    if sweep is not None:
        line["sweep"] = sweep
    del cand, cm_row, cm_rank, losses, scores, z2
    torch.cuda.empty_cache()
    return line


def select_e2e(args):
    """The selection pass through the public API with HOST buffers: losses in -> weights + keep mask out, and
    per-query weights in -> selected ranks + doc ids out (the candidate matrix stays resident in HBM, like the
    dataset object of the reference stays in RAM)."""
    from pacednegatives_b200 import weights as W
    from pacednegatives_b200.dataloader import CandidateMatrix, ROW
    dev = torch.device("cuda", LOCAL_RANK)
    Q, C, Mw = min(args.select_queries, 1 << 20), 1000, 1 << 24
    cand = torch.randint(0, 8841823, (Q, C), dtype=torch.int32, device=dev)
    cm = CandidateMatrix(cand, ROW)
    loss_h = torch.empty(Mw).exponential_(1 / 0.7).pin_memory()
    wq_h = torch.rand(Q).pin_memory()
    w_h = torch.empty(Mw).pin_memory()
    keep_h = torch.empty(Mw, dtype=torch.uint8).pin_memory()
    idx_h = torch.empty(Q, dtype=torch.int32).pin_memory()
    doc_h = torch.empty(Q, dtype=torch.int32).pin_memory()
    eta = torch.tensor([0.3466], device=dev)

    def step(i):
        w, keep, _ = W.spl_weights(loss_h.cuda(non_blocking=True), eta, 0.0, 0, want_keep=True, want_stats=True)
        w_h.copy_(w, non_blocking=True)
        keep_h.copy_(keep, non_blocking=True)
        idx, doc = cm.select_items(wq_h.cuda(non_blocking=True), check=False)
        idx_h.copy_(idx, non_blocking=True)
        doc_h.copy_(doc, non_blocking=True)
        torch.cuda.current_stream().synchronize()
    for i in range(3):
        step(i)
    ms = timed_region(None, step, max(3, min(args.steps, 10)))
    n = max(3, min(args.steps, 10))
    return {"value": Q * n / (ms / 1e3), "unit": "queries/s",
            "h2d_bytes_per_step": int(Mw * 4 + Q * 4), "d2h_bytes_per_step": int(Mw * 5 + Q * 8),
            "ms_per_step": ms / n, "what": f"{Mw} losses -> weights + keep mask, {Q} per-query weights -> rank + doc id"}


# ------------------------------------------------------------------------------------------------
# workload: the LCE training step (BASELINE.json configs[2])
def run_train(args, dist):
    from pacednegatives_b200 import _lib
    from pacednegatives_b200.lceT5 import LCEModel
    from pacednegatives_b200.scorer import T5_BASE, T5_LARGE, random_state_dict, flops_per_doc
    lib = _lib.lib()
    large = args.model == "large"
    S, T = 512, 2
    k = args.negatives if args.negatives else (31 if large else 15)
    Q = args.train_queries
    dims = T5_LARGE if large else T5_BASE
    # T5-large at 1+31 x 512: 512 documents of saved activations (~130 GB) do not fit beside the weights; walk the
    # batch in micro-batches of 8 queries (256 documents) by default
    mq = args.micro_queries if args.micro_queries >= 0 else (8 if large else 0)
    dr = float(getattr(args, "train_dropout", 0.0) or 0.0)
    hp = dict(n=k, use_mean=True, ignore_index=-100, lr=1e-3, meta_lr=1e-3, warmup_steps=0,
              total_steps=1000000, eta=0.3466, micro_queries=mq, dropout_rate=dr, seed=1)
    tag = "T5-large" if large else "monoT5-base"
    cfg_i = 3 if large else 2
    opts = dict(kv.split("=") for kv in (args.set_option or []))
    m = LCEModel(hp, dims=dims, state_dict=random_state_dict(dims, seed=0), cta_group=args.cta_group, options=opts)
    total = args.steps + args.warmup
    nbuf = min(total, 4)
    plab = torch.tensor([[1176, 1]] * Q, dtype=torch.int64)
    nlab = torch.tensor([[6136, 1]] * (Q * k), dtype=torch.int64)
    host = []
    for i in range(nbuf):
        ids = torch.from_numpy(synth_ids(Q * (1 + k), S, 4567 + 1000 * RANK + i))
        host.append(({"input_ids": ids[:Q].contiguous().pin_memory(), "labels": plab.pin_memory()},
                     {"input_ids": ids[Q:].contiguous().pin_memory(), "labels": nlab.pin_memory()}))
    dev = [tuple({kk: v.cuda() for kk, v in part.items()} for part in b) for b in host]
    loss_host = torch.empty(2, dtype=torch.float32).pin_memory()

    def step_dev(i):
        m.training_step(dev[i % nbuf], i)

    def step_e2e(i):
        b = tuple({kk: v.cuda(non_blocking=True) for kk, v in part.items()} for part in host[i % nbuf])
        out = m.training_step(b, i)
        loss_host[0].copy_(out["main_loss"].detach(), non_blocking=True)
        loss_host[1].copy_(out["meta_loss"].detach(), non_blocking=True)
        torch.cuda.current_stream().synchronize()   # the caller logs the step's losses

    for i in range(args.warmup):
        step_dev(i)
    clocks = ClockSampler(LOCAL_RANK)
    clocks.start()
    lib.pn_prof_enable(0)
    l0 = _lib.launch_count()
    ms = timed_region(dist, step_dev, args.steps)
    launches = _lib.launch_count() - l0
    clk = clocks.stop()
    psteps = max(1, min(args.steps, 3))
    lib.pn_prof_enable(1)
    ms_prof = timed_region(dist, step_dev, psteps)
    lib.pn_prof_enable(0)
    prof = {name: prof_read(kind) for name, kind in (("gemm", 0), ("attn", 1), ("wgrad", 3), ("attn_bwd", 4))}
    h_ms, h_n, h_bytes = prof_read(5)
    # exposed gradient exchange: the same steps with the all-reduce switched off (each rank keeps its own gradient)
    ms_noreduce = None
    if dist is not None:
        m.model.skip_grad_reduce = True
        for i in range(2):
            step_dev(i)
        ms_noreduce = timed_region(dist, step_dev, args.steps)
        m.model.skip_grad_reduce = False
    for i in range(2):
        step_e2e(i)
    e2e_ms = timed_region(dist, step_e2e, args.steps)

    queries = Q * args.steps * WORLD
    value = queries / (ms / 1e3)
    e2e_value = queries / (e2e_ms / 1e3)
    peaks = measured_peaks()
    # forward + backward (2x forward) with the redundant no-grad forward elided: 3x the forward count; with dropout
    # the no-grad forward sees other masks and is run: 4x
    step_flops = Q * (1 + k) * flops_per_doc(dims, S, T, True) * (4.0 if dr > 0.0 else 3.0)
    g_ms, g_n, g_flops = prof["gemm"]
    gemm_tflops = g_flops / (g_ms / 1e3) / 1e12 if g_ms > 0 else 0.0

    def tf(name):
        t_ms, n, work = prof[name]
        return {"tflops": work / (t_ms / 1e3) / 1e12 if t_ms else 0.0, "launches": int(n),
                "share_of_step": t_ms / ms_prof if ms_prof else None}
    hbm = hbm_gemm_entry(h_ms, h_n, h_bytes, ms_prof, peaks)
    n_params = int(m.model.grads.numel())
    model_tflops = step_flops / (ms / args.steps / 1e3) / 1e12
    line = {
        "metric": f"LCE training queries/s (1+{k} docs per query)", "value": value, "unit": "queries/s",
        "n_gpus": WORLD, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": f"synthetic (random-init {tag} dims, ids U{{2..32099}}, labels true/false + </s>)",
        "config": {"workload": f"LCE training step, {tag}, 1+{k} docs per query, seq 512 (BASELINE configs[{cfg_i}])",
                   "queries_per_step_per_gpu": Q, "negatives": k, "seq_len": S, "label_tokens": T,
                   "micro_queries": mq if 0 < mq < Q else None,
                   "dropout_rate": dr, "options": opts or None,
                   "no_grad_forward": ("elided (identical arithmetic at dropout 0); FLOP count uses 3x forward" if dr == 0.0
                                       else "run in train mode with its own masks; FLOP count uses 4x forward"),
                   "optimizer": "AdamW (transformers rule) on flat fp32 masters + eta meta step",
                   "l2": "activations per step (tens of GB) exceed the 126 MB L2; no flush needed",
                   "parallelism": f"dp{WORLD}: fp32 SUM all-reduce of the flat gradient buffer in 6 ranges overlapped "
                                  "with the backward pass (NCCL), 1/world folded into AdamW; eta gradient all-reduced",
                   "allreduce_bytes_per_step": n_params * 4 if dist is not None else 0,
                   "allreduce_dtype": "fp32",
                   "allreduce_exposed_ms": (ms - ms_noreduce) / args.steps if ms_noreduce is not None else 0.0,
                   "ms_per_step_without_allreduce": ms_noreduce / args.steps if ms_noreduce is not None else None},
        "gpu_launches": int(launches),
        "model_tflops": model_tflops,
        "model_frac_of_sustained_peak": model_tflops / peaks["bf16_sustained"],
        "e2e": {"value": e2e_value, "unit": "queries/s",
                "h2d_bytes_per_step": int(Q * (1 + k) * S * 4 + Q * (1 + k) * T * 8),
                "d2h_bytes_per_step": 8, "ms_per_step": e2e_ms / args.steps},
        "roofline": {"bound": "tensor", "kernel": "gemm_bf16_kernel (tcgen05: tensor-bound forward linears + data gradients)",
                     "achieved": gemm_tflops, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
                     "frac": gemm_tflops / peaks["bf16_sustained"], "peak_source": peaks["src"] + " sustained",
                     "launches": int(g_n), "avg_launch_ms": g_ms / max(g_n, 1),
                     "share_of_step": g_ms / ms_prof if ms_prof else None, "traffic": None,
                     "model_tflops": model_tflops, "model_frac": model_tflops / peaks["bf16_sustained"],
                     "wgrad_tflops": tf("wgrad")["tflops"], "wgrad_share": tf("wgrad")["share_of_step"],
                     "attn_tflops": tf("attn")["tflops"], "attn_share": tf("attn")["share_of_step"],
                     "attn_bwd_tflops": tf("attn_bwd")["tflops"], "attn_bwd_share": tf("attn_bwd")["share_of_step"],
                     "hbm_gemm_gbs": hbm["achieved"], "hbm_gemm_frac": hbm["frac"],
                     "hbm_gemm_share": hbm["share_of_step"]},
        "clocks": clk,
    }
    del m, dev
    torch.cuda.empty_cache()
    return line


def train_flat(tr):
    """The training line as flat scalars for the re-rank line's `config` (the driver keeps every key of `config`)."""
    c, r = tr["config"], tr["roofline"]
    out = {"train_workload": c["workload"], "train_queries_per_s": tr["value"], "train_ms_per_step": tr["ms_per_step"],
           "train_e2e_queries_per_s": tr["e2e"]["value"], "train_e2e_ms_per_step": tr["e2e"]["ms_per_step"],
           "train_queries_per_step_per_gpu": c["queries_per_step_per_gpu"], "train_negatives": c["negatives"],
           "train_model_tflops": tr["model_tflops"], "train_model_frac": tr["model_frac_of_sustained_peak"],
           "train_flop_count": "3x forward (no-grad forward elided at dropout 0)" if not c.get("dropout_rate") else "4x forward",
           "train_dropout_rate": c.get("dropout_rate", 0.0),
           "train_gpu_launches_per_step": tr["gpu_launches"] / max(tr["steps"], 1),
           "train_allreduce_bytes": c["allreduce_bytes_per_step"], "train_allreduce_dtype": c["allreduce_dtype"],
           "train_allreduce_exposed_ms": c["allreduce_exposed_ms"],
           "train_ms_per_step_without_allreduce": c["ms_per_step_without_allreduce"],
           "train_gemm_tflops": r["achieved"], "train_gemm_frac": r["frac"], "train_gemm_share": r["share_of_step"],
           "train_wgrad_tflops": r["wgrad_tflops"], "train_wgrad_share": r["wgrad_share"],
           "train_attn_tflops": r["attn_tflops"], "train_attn_share": r["attn_share"],
           "train_attn_bwd_tflops": r["attn_bwd_tflops"], "train_attn_bwd_share": r["attn_bwd_share"],
           "train_hbm_gemm_gbs": r["hbm_gemm_gbs"], "train_hbm_gemm_share": r["hbm_gemm_share"],
           "train_sm_mhz": (tr.get("clocks") or {}).get("sm_mhz")}
    if tr.get("cpu_baseline"):
        out["train_cpu_baseline_queries_per_s"] = tr["cpu_baseline"]["value"]
        out["train_cpu_baseline_sample"] = tr["cpu_baseline"]["sample"]
    return out


def select_flat(sel):
    out = {"select_queries_per_s": sel["value"], "select_ms_per_pass": sel["ms_per_step"],
           "select_queries": sel["config"]["queries"], "select_weight_elems": sel["config"]["weight_elems"]}
    short = ["weights", "table_update", "level_rank_major", "level_row_gather", "binomial", "lce_restricted", "topk_metrics"]
    for name, k in zip(short, sel["kernels"]):
        out[f"select_{name}_gbs"] = k["achieved_gbs"]
        out[f"select_{name}_frac"] = k["frac_of_hbm_peak"]
        out[f"select_{name}_ms"] = k["ms"]
    return out


# ------------------------------------------------------------------------------------------------
# --impl reference: the reference's own CPU code path for the same metric, on rank 0's host cores
def reference_rerank(args):
    S = 512
    n = args.ref_docs_per_step
    ids = torch.from_numpy(synth_ids(n, S, 2345).astype(np.int64))
    run1, how = cpu_rerank_fn(n, S)
    from oracle import metrics as om
    for _ in range(args.warmup):
        run1(ids)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        s = run1(ids)
        sc = np.asarray(s, dtype=np.float32).reshape(1, -1)
        om.batch_metrics(sc, np.arange(n, dtype=np.int32).reshape(1, -1),
                         np.zeros((1, n), dtype=np.uint8), min(10, n))
    dt = time.perf_counter() - t0
    v = n * args.steps / dt
    return {"impl": "reference", "metric": "monoT5-base rerank pairs/s", "value": v, "unit": "pairs/s",
            "n_gpus": WORLD, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "monoT5-base re-rank, BM25 top-1000, seq 512 (BASELINE configs[1])",
                       "sample_docs_per_step": n, "seq_len": S},
            "gpu_launches": 0,
            "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": torch.get_num_threads(),
                             "host_cpus": os.cpu_count(), "kind": "port",
                             "sample": f"{n} docs x seq {S} per step, {how}"},
            "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def reference_train(args, steps=None, warmup=None):
    large = args.model == "large"
    S, k = 512, args.negatives if args.negatives else (31 if large else 15)
    tag, cfg_i = ("T5-large", 3) if large else ("monoT5-base", 2)
    n = args.ref_train_docs
    steps = args.steps if steps is None else steps
    warmup = args.warmup if warmup is None else warmup
    v, how, dt = hf_train_sample(n, S, steps, warmup, large=large)
    qs = v / (1 + k)
    return {"impl": "reference", "metric": f"LCE training queries/s (1+{k} docs per query)", "value": qs,
            "unit": "queries/s", "n_gpus": WORLD, "steps": steps, "warmup": warmup,
            "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"LCE training step, {tag}, 1+{k} docs per query, seq 512 (BASELINE configs[{cfg_i}])",
                       "sample_docs_per_step": n, "seq_len": S},
            "gpu_launches": 0,
            "cpu_baseline": {"value": qs, "unit": "queries/s", "cores": torch.get_num_threads(),
                             "host_cpus": os.cpu_count(), "kind": "port",
                             "sample": f"{n} docs x seq {S} per step x {steps} steps, {how}; docs/s / {1 + k}"},
            "e2e": {"value": qs, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def reference_select(args):
    b = cpu_select_baseline()
    v = b["queries_per_s"]
    return {"impl": "reference", "metric": "paced selection + weighting kernels, queries/s through one pass of all kernels",
            "value": v, "unit": "queries/s", "n_gpus": WORLD, "steps": 1, "warmup": 0, "ms_per_step": 1e3 / v,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32/f64", "data": "synthetic",
            "config": {"workload": "paced negative sampling + weighting kernel sweep (BASELINE configs[4])",
                       "sample": "LCEWeights loop M in {64,1024,16384}, LCEDataset.sample x1000, get_items x100000, "
                                 "trec_eval restatement 200 queries", **{f"ref_{k_}": v_ for k_, v_ in b.items()}},
            "gpu_launches": 0,
            "cpu_baseline": {"value": v, "unit": "queries/s", "cores": 1, "host_cpus": os.cpu_count(), "kind": "port",
                             "sample": "reference-shaped Python loops (oracle/baseline.py), one core"},
            "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def main_reference(args):
    if RANK != 0:
        return
    claim_stdout()
    torch.set_num_threads(os.cpu_count() or 1)
    if args.workload == "rerank":
        return emit(reference_rerank(args))
    if args.workload == "train":
        return emit(reference_train(args))
    if args.workload == "select":
        return emit(reference_select(args))
    # "all": the line is the re-rank half (same metric / config as the b200 arm's line); the training half on a
    # bounded sample (at most 3 timed steps of `--ref-train-docs` documents) as flat scalars in `config`
    line = reference_rerank(args)
    try:
        tr = reference_train(args, steps=max(1, min(args.steps, 3)), warmup=1)
        line["config"].update({"train_workload": tr["config"]["workload"], "train_queries_per_s": tr["value"],
                               "train_ms_per_step": tr["ms_per_step"],
                               "train_sample": tr["cpu_baseline"]["sample"]})
        print(json.dumps(tr), file=sys.stderr, flush=True)
    except Exception as e:
        line["config"]["train_error"] = f"{type(e).__name__}: {e}"[:200]
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="all", choices=["all", "rerank", "train", "select"])
    ap.add_argument("--queries-per-step", type=int, default=4)
    ap.add_argument("--chunk-docs", type=int, default=1000)
    ap.add_argument("--cta-group", type=int, default=2)
    ap.add_argument("--cpu-sample-docs", type=int, default=32)
    ap.add_argument("--ref-docs-per-step", type=int, default=16)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-library-baseline", action="store_true")
    ap.add_argument("--no-select", action="store_true", help="workload all: skip the short selection-kernel pass")
    ap.add_argument("--sweep", action="store_true", help="workload select: also run the SURVEY 8d cfg5 sweep")
    ap.add_argument("--select-queries", type=int, default=1000000)
    ap.add_argument("--select-elems", type=int, default=1 << 28)
    ap.add_argument("--train-queries", type=int, default=16)
    ap.add_argument("--model", default="base", choices=["base", "large"],
                    help="train workload: base = BASELINE configs[2] (1+15), large = configs[3] (1+31, micro-batched)")
    ap.add_argument("--negatives", type=int, default=0, help="train workload: negatives per query (0 = config default)")
    ap.add_argument("--micro-queries", type=int, default=-1,
                    help="train workload: queries per micro-batch (0 = whole batch at once, -1 = config default)")
    ap.add_argument("--ref-train-docs", type=int, default=4)
    ap.add_argument("--set-option", action="append", metavar="NAME=VALUE",
                    help="train workload: pn_t5_set_option on the model (A/B runs: dgrad_bf16=0, grad_ranges=1 ...)")
    ap.add_argument("--train-dropout", type=float, default=0.0,
                    help="train workload: dropout_rate (0.1 = the reference's train-mode regime: 4 forwards' worth of work)")
    ap.add_argument("--no-dropout-regime", action="store_true",
                    help="workload all: skip the short second training measurement at dropout 0.1")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3   # timing rule: at least 3 warm-up steps
    if args.gpus != WORLD and WORLD == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29511", os.path.abspath(__file__)] + sys.argv[1:]
        os.execv(sys.executable, cmd)
    if args.impl == "reference":
        return main_reference(args)

    from pacednegatives_b200 import _lib
    _lib.require_gpu()
    torch.cuda.set_device(LOCAL_RANK)
    dist = dist_setup(True)
    solo = RANK == 0 and WORLD == 1
    if args.workload == "rerank" or args.workload == "all":
        line = run_rerank(args, dist)
        if solo and not args.no_cpu_baseline:
            v, how, dt = cpu_rerank_sample(args.cpu_sample_docs, 512)
            line["cpu_baseline"] = {"value": v, "unit": "pairs/s", "cores": torch.get_num_threads(),
                                    "host_cpus": os.cpu_count(), "kind": "port",
                                    "sample": f"{args.cpu_sample_docs} docs x seq 512, t5-base, {how}, {dt:.1f} s"}
    if args.workload == "train" or args.workload == "all":
        tr = run_train(args, dist)
        if solo and not args.no_cpu_baseline:
            try:
                v, how, dt = hf_train_sample(4, 512, 1, 1, large=args.model == "large")
                kk = tr["config"]["negatives"]
                tr["cpu_baseline"] = {"value": v / (1 + kk), "unit": "queries/s", "cores": torch.get_num_threads(),
                                      "host_cpus": os.cpu_count(), "kind": "port",
                                      "sample": f"4 docs x seq 512 x 1 step ({dt:.1f} s), {how}; docs/s / {1 + kk}"}
            except Exception as e:   # transformers missing: say so rather than invent a number
                tr["cpu_baseline"] = {"value": None, "unit": "queries/s", "kind": "port", "sample": f"failed: {e}"}
        if args.workload == "train":
            line = tr
        else:
            line["config"].update(train_flat(tr))
            if RANK == 0:
                print(json.dumps(tr), file=sys.stderr, flush=True)
            if not args.no_dropout_regime and args.train_dropout == 0.0:
                # the reference's own regime (Lightning train mode: dropout 0.1 in both forwards), a short run
                import copy
                a2 = copy.copy(args)
                a2.train_dropout, a2.steps = 0.1, max(3, min(args.steps, 5))
                try:
                    td = run_train(a2, dist)
                    line["config"].update({"train_dropout01_queries_per_s": td["value"],
                                           "train_dropout01_ms_per_step": td["ms_per_step"],
                                           "train_dropout01_model_frac": td["model_frac_of_sustained_peak"],
                                           "train_dropout01_flop_count": "4x forward",
                                           "train_dropout01_steps": td["steps"]})
                    if RANK == 0:
                        print(json.dumps(td), file=sys.stderr, flush=True)
                except Exception as e:
                    line["config"]["train_dropout01_error"] = f"{type(e).__name__}: {e}"[:200]
    if args.workload == "select":
        line = run_select(args, dist)
        if solo:
            line["e2e"] = select_e2e(args)
            if not args.no_cpu_baseline:
                b = cpu_select_baseline()
                line["cpu_baseline"] = {"value": b["queries_per_s"], "unit": "queries/s", "cores": 1,
                                        "host_cpus": os.cpu_count(), "kind": "port",
                                        "sample": "reference-shaped Python loops (oracle/baseline.py): LCEWeights.forward "
                                                  "M in {64,1024,16384}, LCEDataset.sample x1000, get_items x100000, "
                                                  "trec_eval restatement x200",
                                        **{k_: v_ for k_, v_ in b.items() if k_ != "queries_per_s"}}
    if args.workload == "all" and not args.no_select:
        try:
            sel = run_select(args, dist, quick=True)
            line["config"].update(select_flat(sel))
            if RANK == 0:
                print(json.dumps(sel), file=sys.stderr, flush=True)
        except Exception as e:
            line["config"]["select_error"] = f"{type(e).__name__}: {e}"[:200]
    if solo and args.workload in ("all", "rerank") and not args.no_library_baseline:
        line["config"].update(library_baseline())
    if RANK == 0:
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

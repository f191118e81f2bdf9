#!/usr/bin/env python
"""bench.py — the reference's headline metrics on B200 (BASELINE.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload rerank|train|select] [--impl reference]

workload "rerank" (default; BASELINE.json configs[1]): monoT5-base re-rank of synthetic BM25 top-1000
lists, seq 512, bf16 tensor-core GEMMs with fp32 accumulation.  One step = `--queries-per-step` queries
x 1000 candidate prompts scored (encoder + one decoder step + two-row head), ranked, and evaluated
(top-10, nDCG@10, RR) — SURVEY §8 a1 + a9 + a12.  metric = pairs/s over all ranks.

workload "train" (configs[2]): one LCE training step per rank (Q queries x (1 + k) documents).
workload "select" (configs[4]): the paced-selection + weighting bandwidth kernels.

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
# workload: re-rank (BASELINE.json configs[1])
def cpu_rerank_sample(n_docs, S, threads=None):
    """The reference's CPU path on a bounded sample: HuggingFace T5ForConditionalGeneration fp32 forward
    (what lceT5.py:132 / pyterrier_t5 call) when transformers is importable, else the oracle's torch-CPU
    restatement; then the monoT5 score formula (distill/wrapper.py:34-38).  Returns (pairs/s, how)."""
    if threads:
        torch.set_num_threads(threads)
    ids = torch.from_numpy(synth_ids(n_docs, S, 999).astype(np.int64))
    how = None
    try:
        from transformers import T5Config, T5ForConditionalGeneration
        cfg = T5Config(vocab_size=32128, d_model=768, d_kv=64, d_ff=3072, num_layers=12,
                       num_decoder_layers=12, num_heads=12, dropout_rate=0.0,
                       decoder_start_token_id=0, pad_token_id=0, eos_token_id=1)
        torch.manual_seed(0)
        model = T5ForConditionalGeneration(cfg).eval()
        dec = torch.zeros(n_docs, 1, dtype=torch.long)

        def run(x):
            with torch.no_grad():
                lg = model(input_ids=x, decoder_input_ids=dec[: x.shape[0]]).logits
                return torch.log_softmax(lg[:, 0, [1176, 6136]], dim=1)[:, 0]
        how = "transformers T5ForConditionalGeneration fp32 eager CPU"
    except Exception:  # transformers missing or incompatible: the oracle's restatement
        from oracle import t5 as ot
        w = ot.init_weights(ot.T5_BASE, seed=0)

        def run(x):
            return ot.score(w, ot.T5_BASE, x.numpy(), 1176, 6136)[0]
        how = "oracle.t5 torch-CPU fp32 restatement"
    bs = 8
    run(ids[:bs])  # warm-up
    t0 = time.perf_counter()
    for lo in range(0, n_docs, bs):
        run(ids[lo:lo + bs])
    dt = time.perf_counter() - t0
    return n_docs / dt, how, dt


def bench_rerank(args):
    from pacednegatives_b200 import _lib
    from pacednegatives_b200.scorer import MonoT5Scorer, T5_BASE, random_state_dict, flops_per_doc
    _lib.require_gpu()
    torch.cuda.set_device(LOCAL_RANK)
    dist = dist_setup(True)
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

    def step_dev(i):
        scorer.rerank(dev_ids[i % nbuf].view(Qs, C, S), docids_d, rels_d, K)

    def step_e2e(i):
        ids = host_ids[i % nbuf].cuda(non_blocking=True)
        out = scorer.rerank(ids.view(Qs, C, S), docids_h.cuda(non_blocking=True),
                            rels_h.cuda(non_blocking=True), K)
        for k, v in out_host.items():
            v.copy_(out[k], non_blocking=True)
        torch.cuda.current_stream().synchronize()   # the caller reads the step's ranking and metrics

    for i in range(args.warmup):
        step_dev(i)
    clocks = ClockSampler(LOCAL_RANK)
    clocks.start()
    lib.pn_prof_enable(1)
    l0 = _lib.launch_count()
    ms = timed_region(dist, step_dev, args.steps)
    launches = _lib.launch_count() - l0
    lib.pn_prof_enable(0)
    clk = clocks.stop()
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
    line = {
        "metric": "monoT5-base rerank pairs/s", "value": value, "unit": "pairs/s", "n_gpus": WORLD,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic (random-init t5-base dims, ids U{2..32099}, graded qrels)",
        "config": {"workload": "monoT5-base re-rank, BM25 top-1000, seq 512 (BASELINE configs[1])",
                   "queries_per_step_per_gpu": Qs, "candidates": C, "seq_len": S,
                   "chunk_docs": args.chunk_docs, "cta_group": args.cta_group,
                   "l2": "inputs+activations per step (>1 GB) exceed the 126 MB L2; no flush needed",
                   "parallelism": f"queries sharded over {WORLD} rank(s), no data-path collective"},
        "gpu_launches": int(launches),
        "model_tflops": value * fpd / 1e12 / WORLD,
        "model_frac_of_sustained_peak": value * fpd / 1e12 / WORLD / peaks["bf16_sustained"],
        "e2e": {"value": e2e_value, "unit": "pairs/s",
                "h2d_bytes_per_step": int(host_ids[0].numel() * 4 + rels_h.numel() + docids_h.numel() * 4),
                "d2h_bytes_per_step": int(sum(v.numel() * v.element_size() for v in out_host.values())),
                "ms_per_step": e2e_ms / args.steps},
        "roofline": {"bound": "tensor", "kernel": "gemm_bf16_kernel (tcgen05; tensor-bound T5 linears, HBM-bound launches listed apart)",
                     "achieved": gemm_tflops, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
                     "frac": gemm_tflops / peaks["bf16_sustained"], "peak_source": peaks["src"] + " sustained",
                     "launches": int(g_n), "avg_launch_ms": g_ms / max(g_n, 1),
                     "share_of_step": g_ms / ms if ms else None,
                     "attention": {"tflops": a_flops / (a_ms / 1e3) / 1e12 if a_ms else 0.0,
                                   "launches": int(a_n), "share_of_step": a_ms / ms if ms else None},
                     "hbm_bound_gemms": hbm_gemm_entry(h_ms, h_n, h_bytes, ms, peaks),
                     "traffic": traffic, "traffic_source": traffic_src},
        "clocks": clk,
    }
    if RANK == 0:
        if WORLD == 1 and not args.no_cpu_baseline:
            v, how, dt = cpu_rerank_sample(args.cpu_sample_docs, S)
            line["cpu_baseline"] = {"value": v, "unit": "pairs/s", "cores": torch.get_num_threads(),
                                    "host_cpus": os.cpu_count(), "kind": "port",
                                    "sample": f"{args.cpu_sample_docs} docs x seq {S}, t5-base, {how}, {dt:.1f} s"}
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


# workload: paced selection + weighting bandwidth kernels (BASELINE.json configs[4])
def bench_select(args):
    from pacednegatives_b200 import _lib
    from pacednegatives_b200 import weights as W, loss as L, metrics as M
    from pacednegatives_b200.dataloader import CandidateMatrix, ROW, RANK as RANKM
    _lib.require_gpu()
    torch.cuda.set_device(LOCAL_RANK)
    dist = dist_setup(True)
    dev = torch.device("cuda", LOCAL_RANK)
    Q, C, k = args.select_queries, 1000, 15
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
    Mw = args.select_elems
    losses = torch.empty(Mw, device=dev).exponential_(1 / 0.7, generator=g)
    eta = torch.tensor([0.3466], device=dev)
    wq = torch.rand(Q, device=dev, generator=g)
    table, _ = cm_row.binomial_table(0.3466, 0.01)
    G = 16
    z2 = torch.randn(Q * G, 2, device=dev, generator=g)
    gq = torch.ones(Q, device=dev)
    Qm = min(Q, 200000)
    scores = torch.randn(Qm, C, device=dev, generator=g)
    rels = (torch.rand(Qm, C, device=dev, generator=g) > 0.95).to(torch.uint8)
    docids = cand[:Qm]
    row_tab = torch.ones(1 << 24, device=dev)
    ce_p = torch.rand(1 << 24, device=dev, generator=g)
    ce_n = torch.rand(1 << 24, device=dev, generator=g)
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
        for _ in range(max(args.warmup, 3)):
            fn()
        ts = []
        for _ in range(args.steps):
            flush.zero_()     # inputs that fit the 126 MB L2 would otherwise be read from cache
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = sum(ts) / len(ts)
        tot_ms += ms
        gbs = nbytes / (ms / 1e3) / 1e9
        res.append({"kernel": name, "ms": ms, "algorithmic_bytes": nbytes, "achieved_gbs": gbs,
                    "frac_of_hbm_peak": gbs / peaks["hbm_gbs"]})
    launches = _lib.launch_count() - l0
    top = res[0]
    sel_traffic, sel_traffic_src = ncu_traffic("select", Mw)   # dram bytes of one spl_weights launch, scaled by elements
    line = {"metric": "paced selection + weighting kernels, queries/s through one pass of all kernels",
            "value": Q * WORLD / (tot_ms / 1e3), "unit": "queries/s", "n_gpus": WORLD, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": tot_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32/i32/f64", "data": "synthetic (SURVEY cfg5 generators)",
            "config": {"workload": "paced negative sampling + weighting kernel sweep (BASELINE configs[4])",
                       "queries": Q, "candidates": C, "k": k, "weight_elems": Mw, "l2": "256 MB flush between iterations"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": top["kernel"], "achieved": top["achieved_gbs"], "peak": peaks["hbm_gbs"],
                         "unit": "GB/s", "frac": top["frac_of_hbm_peak"], "peak_source": peaks["src"],
                         "traffic": sel_traffic, "traffic_source": sel_traffic_src},
            "kernels": res}
    if RANK == 0:
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


def bench_rerank_reference(args):
    """`--impl reference`: the reference's CPU code path for the same metric, on rank 0's host cores."""
    if RANK != 0:
        return
    S = 512
    n = args.ref_docs_per_step
    torch.set_num_threads(os.cpu_count() or 1)
    # K + W steps of n docs each
    t_all = []
    how = None
    ids = torch.from_numpy(synth_ids(n, S, 2345).astype(np.int64))
    try:
        from transformers import T5Config, T5ForConditionalGeneration
        cfg = T5Config(vocab_size=32128, d_model=768, d_kv=64, d_ff=3072, num_layers=12,
                       num_decoder_layers=12, num_heads=12, dropout_rate=0.0,
                       decoder_start_token_id=0, pad_token_id=0, eos_token_id=1)
        torch.manual_seed(0)
        model = T5ForConditionalGeneration(cfg).eval()
        dec = torch.zeros(n, 1, dtype=torch.long)

        def run():
            with torch.no_grad():
                lg = model(input_ids=ids, decoder_input_ids=dec).logits
                s = torch.log_softmax(lg[:, 0, [1176, 6136]], dim=1)[:, 0]
            return s
        how = "transformers T5ForConditionalGeneration fp32 eager CPU + monoT5 score (distill/wrapper.py:34-38)"
    except Exception:
        from oracle import t5 as ot
        w = ot.init_weights(ot.T5_BASE, seed=0)

        def run():
            return ot.score(w, ot.T5_BASE, ids.numpy(), 1176, 6136)[0]
        how = "oracle.t5 torch-CPU fp32 restatement"
    from oracle import metrics as om
    for _ in range(args.warmup):
        run()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        s = run()
        sc = np.asarray(s, dtype=np.float32).reshape(1, -1)
        om.batch_metrics(sc, np.arange(n, dtype=np.int32).reshape(1, -1),
                         np.zeros((1, n), dtype=np.uint8), min(10, n))
    dt = time.perf_counter() - t0
    v = n * args.steps / dt
    line = {"impl": "reference", "metric": "monoT5-base rerank pairs/s", "value": v, "unit": "pairs/s",
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
    emit(line)


# ------------------------------------------------------------------------------------------------
# workload: the LCE training step (BASELINE.json configs[2])
def hf_train_sample(n_docs, S, steps, warmup, threads=None, large=False):
    """The reference's CPU path for one training step's model work on a bounded sample: HF
    T5ForConditionalGeneration fp32 forward + backward + AdamW over n_docs prompts with 2 label tokens and the
    reference's loss (lceT5.py:149-156).  Returns (docs/s, description, seconds)."""
    if threads:
        torch.set_num_threads(threads)
    from transformers import T5Config, T5ForConditionalGeneration
    d, ff, L, H = (1024, 4096, 24, 16) if large else (768, 3072, 12, 12)
    cfg = T5Config(vocab_size=32128, d_model=d, d_kv=64, d_ff=ff, num_layers=L, num_decoder_layers=L,
                   num_heads=H, dropout_rate=0.0, decoder_start_token_id=0, pad_token_id=0, eos_token_id=1)
    torch.manual_seed(0)
    model = T5ForConditionalGeneration(cfg).train()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3, eps=1e-6, weight_decay=0.0)
    ids = torch.from_numpy(synth_ids(n_docs, S, 4567).astype(np.int64))
    labels = torch.tensor([[1176, 1]] + [[6136, 1]] * (n_docs - 1))
    loss_fn = torch.nn.CrossEntropyLoss(ignore_index=-100, reduction="none")

    def run():
        logits = model(input_ids=ids, labels=labels).logits
        ce = loss_fn(logits.view(-1, logits.size(-1)), labels.view(-1))
        opt.zero_grad()
        ce.mean().backward()
        opt.step()
        return float(ce.mean())
    for _ in range(warmup):
        run()
    t0 = time.perf_counter()
    for _ in range(steps):
        run()
    dt = time.perf_counter() - t0
    return n_docs * steps / dt, "transformers T5ForConditionalGeneration fp32 eager CPU fwd+bwd+AdamW", dt


def bench_train(args):
    from pacednegatives_b200 import _lib
    from pacednegatives_b200.lceT5 import LCEModel
    from pacednegatives_b200.scorer import T5_BASE, T5_LARGE, random_state_dict, flops_per_doc
    _lib.require_gpu()
    torch.cuda.set_device(LOCAL_RANK)
    dist = dist_setup(True)
    lib = _lib.lib()
    large = args.model == "large"
    S, T = 512, 2
    k = args.negatives if args.negatives else (31 if large else 15)
    Q = args.train_queries
    dims = T5_LARGE if large else T5_BASE
    # T5-large at 1+31 x 512: 512 documents of saved activations (~130 GB) do not fit beside the weights; walk the
    # batch in micro-batches of 8 queries (256 documents) by default
    mq = args.micro_queries if args.micro_queries >= 0 else (8 if large else 0)
    hp = dict(n=k, use_mean=True, ignore_index=-100, lr=1e-3, meta_lr=1e-3, warmup_steps=0,
              total_steps=1000000, eta=0.3466, micro_queries=mq)
    tag = "T5-large" if large else "monoT5-base"
    cfg_i = 3 if large else 2
    m = LCEModel(hp, dims=dims, state_dict=random_state_dict(dims, seed=0), cta_group=args.cta_group)
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
    lib.pn_prof_enable(1)
    l0 = _lib.launch_count()
    ms = timed_region(dist, step_dev, args.steps)
    launches = _lib.launch_count() - l0
    lib.pn_prof_enable(0)
    clk = clocks.stop()
    prof = {name: prof_read(kind) for name, kind in (("gemm", 0), ("attn", 1), ("wgrad", 3), ("attn_bwd", 4))}
    for i in range(2):
        step_e2e(i)
    e2e_ms = timed_region(dist, step_e2e, args.steps)

    queries = Q * args.steps * WORLD
    value = queries / (ms / 1e3)
    e2e_value = queries / (e2e_ms / 1e3)
    peaks = measured_peaks()
    # forward + backward (2x forward) with the redundant no-grad forward elided: 3x the forward count
    step_flops = Q * (1 + k) * flops_per_doc(dims, S, T, True) * 3.0
    g_ms, g_n, g_flops = prof["gemm"]
    gemm_tflops = g_flops / (g_ms / 1e3) / 1e12 if g_ms > 0 else 0.0

    def tf(name):
        t_ms, n, work = prof[name]
        return {"tflops": work / (t_ms / 1e3) / 1e12 if t_ms else 0.0, "launches": int(n),
                "share_of_step": t_ms / ms if ms else None}
    line = {
        "metric": f"LCE training queries/s (1+{k} docs per query)", "value": value, "unit": "queries/s",
        "n_gpus": WORLD, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": f"synthetic (random-init {tag} dims, ids U{{2..32099}}, labels true/false + </s>)",
        "config": {"workload": f"LCE training step, {tag}, 1+{k} docs per query, seq 512 (BASELINE configs[{cfg_i}])",
                   "queries_per_step_per_gpu": Q, "negatives": k, "seq_len": S, "label_tokens": T,
                   "micro_queries": mq if 0 < mq < Q else None,
                   "no_grad_forward": "elided (identical arithmetic at dropout 0); FLOP count uses 3x forward",
                   "optimizer": "AdamW (transformers rule) on flat fp32 masters + eta meta step",
                   "l2": "activations per step (tens of GB) exceed the 126 MB L2; no flush needed",
                   "parallelism": f"dp{WORLD}: one all-reduce of the flat fp32 gradient buffer per step"},
        "gpu_launches": int(launches),
        "model_tflops": step_flops / (ms / args.steps / 1e3) / 1e12,
        "model_frac_of_sustained_peak": step_flops / (ms / args.steps / 1e3) / 1e12 / peaks["bf16_sustained"],
        "e2e": {"value": e2e_value, "unit": "queries/s",
                "h2d_bytes_per_step": int(Q * (1 + k) * S * 4 + Q * (1 + k) * T * 8),
                "d2h_bytes_per_step": 8, "ms_per_step": e2e_ms / args.steps},
        "roofline": {"bound": "tensor", "kernel": "gemm_bf16_kernel (tcgen05: tensor-bound forward linears + data gradients)",
                     "achieved": gemm_tflops, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
                     "frac": gemm_tflops / peaks["bf16_sustained"], "peak_source": peaks["src"] + " sustained",
                     "launches": int(g_n), "avg_launch_ms": g_ms / max(g_n, 1),
                     "share_of_step": g_ms / ms if ms else None,
                     "wgrad": tf("wgrad"), "attention": tf("attn"), "attention_bwd": tf("attn_bwd"),
                     "hbm_bound_gemms": hbm_gemm_entry(*prof_read(5), ms, peaks),
                     "traffic": None},
        "clocks": clk,
    }
    if RANK == 0:
        if WORLD == 1 and not args.no_cpu_baseline:
            try:
                v, how, dt = hf_train_sample(4, S, 1, 1, large=large)
                line["cpu_baseline"] = {"value": v / (1 + k), "unit": "queries/s", "cores": torch.get_num_threads(),
                                        "host_cpus": os.cpu_count(), "kind": "port",
                                        "sample": f"4 docs x seq {S} x 1 step ({dt:.1f} s), {how}; docs/s / {1 + k}"}
            except Exception as e:   # transformers missing: say so rather than invent a number
                line["cpu_baseline"] = {"value": None, "unit": "queries/s", "kind": "port", "sample": f"failed: {e}"}
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


def bench_train_reference(args):
    """`--impl reference --workload train`: HF fp32 forward+backward+AdamW on the host cores, rank 0 only."""
    if RANK != 0:
        return
    large = args.model == "large"
    S, k = 512, args.negatives if args.negatives else (31 if large else 15)
    tag, cfg_i = ("T5-large", 3) if large else ("monoT5-base", 2)
    n = args.ref_train_docs
    torch.set_num_threads(os.cpu_count() or 1)
    v, how, dt = hf_train_sample(n, S, args.steps, args.warmup, large=large)
    qs = v / (1 + k)
    line = {"impl": "reference", "metric": f"LCE training queries/s (1+{k} docs per query)", "value": qs,
            "unit": "queries/s", "n_gpus": WORLD, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"LCE training step, {tag}, 1+{k} docs per query, seq 512 (BASELINE configs[{cfg_i}])",
                       "sample_docs_per_step": n, "seq_len": S},
            "gpu_launches": 0,
            "cpu_baseline": {"value": qs, "unit": "queries/s", "cores": torch.get_num_threads(),
                             "host_cpus": os.cpu_count(), "kind": "port",
                             "sample": f"{n} docs x seq {S} per step, {how}; docs/s / {1 + k}"},
            "e2e": {"value": qs, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="rerank", choices=["rerank", "train", "select"])
    ap.add_argument("--queries-per-step", type=int, default=4)
    ap.add_argument("--chunk-docs", type=int, default=1000)
    ap.add_argument("--cta-group", type=int, default=2)
    ap.add_argument("--cpu-sample-docs", type=int, default=32)
    ap.add_argument("--ref-docs-per-step", type=int, default=16)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--select-queries", type=int, default=1000000)
    ap.add_argument("--select-elems", type=int, default=1 << 28)
    ap.add_argument("--train-queries", type=int, default=16)
    ap.add_argument("--model", default="base", choices=["base", "large"],
                    help="train workload: base = BASELINE configs[2] (1+15), large = configs[3] (1+31, micro-batched)")
    ap.add_argument("--negatives", type=int, default=0, help="train workload: negatives per query (0 = config default)")
    ap.add_argument("--micro-queries", type=int, default=-1,
                    help="train workload: queries per micro-batch (0 = whole batch at once, -1 = config default)")
    ap.add_argument("--ref-train-docs", type=int, default=4)
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3   # timing rule: at least 3 warm-up steps
    if args.gpus != WORLD and WORLD == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29511", os.path.abspath(__file__)] + sys.argv[1:]
        os.execv(sys.executable, cmd)
    if args.impl == "reference":
        if args.workload == "rerank":
            return bench_rerank_reference(args)
        if args.workload == "train":
            return bench_train_reference(args)
        raise SystemExit("reference arm: workload not implemented")
    if args.workload == "rerank":
        return bench_rerank(args)
    if args.workload == "select":
        return bench_select(args)
    if args.workload == "train":
        return bench_train(args)
    raise SystemExit(f"workload {args.workload} not implemented yet")


if __name__ == "__main__":
    main()

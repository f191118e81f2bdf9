"""Time spl_weights at 2^28 elements (CUDA events, 256 MB L2 flush between iterations); PN_W_CTAS picks the grid."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pacednegatives_b200 import _lib, weights as W
_lib.require_gpu()
n = 1 << 28
loss = torch.empty(n, device="cuda").exponential_(1 / 0.7)
eta = torch.tensor([0.3466], device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for keep, stats in ((True, True), (False, False)):
    ts = []
    for it in range(8):
        flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        W.spl_weights(loss, eta, 0.0, 0, want_keep=keep, want_stats=stats)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = min(ts[2:])
    b = n * (9.0 if keep else 8.0)
    print(f"PN_W_CTAS={os.environ.get('PN_W_CTAS','8')} keep={keep} stats={stats} ms={t:.4f} GB/s={b / t / 1e6:.0f}", flush=True)

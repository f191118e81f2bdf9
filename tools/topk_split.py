"""Where the top-k kernel's time goes: winners only (no rels) vs winners + measures, at the bench size."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pacednegatives_b200 import _lib, metrics as M
_lib.require_gpu()
Q, C = 200000, 1000
torch.manual_seed(0)
scores = torch.randn(Q, C, device="cuda")
rels = (torch.rand(Q, C, device="cuda") < 0.05).to(torch.uint8)
docids = torch.arange(Q * C, dtype=torch.int32, device="cuda").view(Q, C)
def t(fn):
    ts = []
    for _ in range(4):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts[1:])
print("winners only (K=10)     ms", round(t(lambda: M.topk_metrics(scores, docids, None, 10)), 3))
print("winners + measures K=10 ms", round(t(lambda: M.topk_metrics(scores, docids, rels, 10)), 3))
print("winners only (K=1)      ms", round(t(lambda: M.topk_metrics(scores, docids, None, 1)), 3))
print("winners + measures K=1  ms", round(t(lambda: M.topk_metrics(scores, docids, rels, 1)), 3))
print("bitonic K=33 + measures ms", round(t(lambda: M.topk_metrics(scores, docids, rels, 33)), 3))

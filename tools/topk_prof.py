"""One launch shape of pn_topk_metrics for ncu captures (200 000 queries x 1 000 candidates, 5 % relevant, K = 10)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pacednegatives_b200 import metrics as M, _lib
_lib.require_gpu()
g = torch.Generator(device='cuda').manual_seed(1)
Q, C = 200000, 1000
scores = torch.randn(Q, C, device='cuda', generator=g)
rels = (torch.rand(Q, C, device='cuda', generator=g) > 0.95).to(torch.uint8)
docids = torch.randint(0, 8841823, (Q, C), device='cuda', dtype=torch.int32, generator=g)
for _ in range(3):
    M.topk_metrics(scores, docids, rels, 10)
torch.cuda.synchronize()
print("ok")

"""Timing of pn_topk_metrics (200 000 queries x 1 000 candidates, K = 10) at three relevance densities and with
position-correlated scores; PN_TOPK_WARP selects the kernel generation (2 default, 1, 0)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pacednegatives_b200 import metrics as M, _lib
_lib.require_gpu()
v = os.environ.get("PN_TOPK_WARP", "2")
for frac, tag in ((0.95, "5 % relevant"), (0.998, "0.2 % relevant"), (0.70, "30 % relevant")):
    g = torch.Generator(device='cuda').manual_seed(1)
    Q, C = 200000, 1000
    scores = torch.randn(Q, C, device='cuda', generator=g)
    rels = (torch.rand(Q, C, device='cuda', generator=g) > frac).to(torch.uint8)
    docids = torch.randint(0, 8841823, (Q, C), device='cuda', dtype=torch.int32, generator=g)
    for _ in range(3): out = M.topk_metrics(scores, docids, rels, 10)
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); out = M.topk_metrics(scores, docids, rels, 10); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sum(ts) / len(ts)
    print(f"PN_TOPK_WARP={v} {tag}: top-k + metrics {Q} x {C}: {ms:.3f} ms = {Q*(C*9+96)/ms/1e6:.0f} GB/s; checksum ndcg "
          f"{out['ndcg'].double().sum().item():.6f} ap {out['ap'].double().sum().item():.6f} rr {out['rr'].double().sum().item():.6f} "
          f"ids {out['top_ids'].long().sum().item()}", flush=True)
# scores falling with the position (a first stage that already ranks well): the tau filter keeps 37 keys
Q, C = 200000, 1000
scores = (-torch.arange(C, device='cuda', dtype=torch.float32)).repeat(Q, 1) + torch.randn(Q, C, device='cuda') * 3.0
rels = (torch.rand(Q, C, device='cuda') > 0.95).to(torch.uint8)
for _ in range(3): out = M.topk_metrics(scores, None, rels, 10)
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(5): out = M.topk_metrics(scores, None, rels, 10)
e1.record(); torch.cuda.synchronize()
print(f"PN_TOPK_WARP={v} position-correlated scores, no docids: {e0.elapsed_time(e1)/5:.3f} ms", flush=True)

#!/bin/bash
TAG=${1:-r4k}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_small_kernels.py tests/test_gpu_t5.py -q -m gpu -k "topk or rerank or empty or argument" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${TAG}_pytest.log | cut -c1-300
for v in 2 1 0; do
PN_TOPK_WARP=$v timeout 120 python - <<PY
import torch, sys
sys.path.insert(0, '.')
from pacednegatives_b200 import metrics as M, _lib
_lib.require_gpu()
g = torch.Generator(device='cuda').manual_seed(1)
Q, C = 200000, 1000
scores = torch.randn(Q, C, device='cuda', generator=g)
rels = (torch.rand(Q, C, device='cuda', generator=g) > 0.95).to(torch.uint8)
docids = torch.randint(0, 8841823, (Q, C), device='cuda', dtype=torch.int32, generator=g)
for _ in range(3): out = M.topk_metrics(scores, docids, rels, 10)
ts = []
for _ in range(5):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record(); out = M.topk_metrics(scores, docids, rels, 10); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = sum(ts) / len(ts)
print(f"PN_TOPK_WARP=$v top-k + metrics {Q} x {C}: {ms:.3f} ms = {Q*(C*9+96)/ms/1e6:.0f} GB/s; checksum ndcg {out['ndcg'].double().sum().item():.6f} ap {out['ap'].double().sum().item():.6f} ids {out['top_ids'].long().sum().item()}")
PY
done

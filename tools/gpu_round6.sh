#!/bin/bash
# round-2 session 9: second-generation warp-per-query top-k kernel — parity, A/B against the first generation,
# ncu --set full; then the round-end sequence (full GPU suite, smoke, default bench, reference arm)
TAG=${1:-r5a}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_small_kernels.py tests/test_gpu_t5.py -q -m gpu -k "topk or rerank or empty or argument" > gpurun_out/${TAG}_pytest_topk.log 2>&1; echo "topk pytest rc=$?"; tail -4 gpurun_out/${TAG}_pytest_topk.log | cut -c1-300
for v in 2 1 2 1; do PN_TOPK_WARP=$v timeout 200 python tools/topk_time.py; done 2>&1 | tee gpurun_out/${TAG}_topk_ab.txt
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
head -c 1200 gpurun_out/${TAG}_bench.json; echo
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "reference rc=$?"
head -c 500 gpurun_out/${TAG}_bench_reference.json; echo
cat > /tmp/topk_prof.py <<'PY'
import sys, torch
sys.path.insert(0, '.')
from pacednegatives_b200 import metrics as M, _lib
_lib.require_gpu()
g = torch.Generator(device='cuda').manual_seed(1)
Q, C = 200000, 1000
scores = torch.randn(Q, C, device='cuda', generator=g)
rels = (torch.rand(Q, C, device='cuda', generator=g) > 0.95).to(torch.uint8)
docids = torch.randint(0, 8841823, (Q, C), device='cuda', dtype=torch.int32, generator=g)
for _ in range(3): M.topk_metrics(scores, docids, rels, 10)
torch.cuda.synchronize()
PY
timeout 100 python /tmp/topk_prof.py && timeout 400 ncu --set full --clock-control none --import-source on -k "regex:topk_warp2_kernel" -s 2 -c 1 -f -o gpurun_out/${TAG}_topk python /tmp/topk_prof.py > gpurun_out/${TAG}_ncu_full_topk.log 2>&1
echo "full topk rc=$?"
ncu -i gpurun_out/${TAG}_topk.ncu-rep --page raw --csv > gpurun_out/${TAG}_topk_raw.csv 2>/dev/null
ls -la gpurun_out | grep ${TAG}

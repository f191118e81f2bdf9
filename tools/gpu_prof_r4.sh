#!/bin/bash
# Round-2 profiles: ncu launch lists of one re-rank step and one training step, then ncu --set full of the forward's
# top kernels and of the warp-per-query top-k kernel.  Each command is first run WITHOUT ncu (must exit 0).
TAG=${1:-r4n}
mkdir -p gpurun_out
R="python bench.py --workload rerank --steps 1 --warmup 3 --queries-per-step 1 --no-cpu-baseline --no-library-baseline"
T="python bench.py --workload train --steps 1 --warmup 3 --no-cpu-baseline"
timeout 200 $R > gpurun_out/${TAG}_small_rerank.json 2>/dev/null && \
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 600 --csv --log-file gpurun_out/${TAG}_rerank_launches.csv $R > gpurun_out/${TAG}_ncu_rerank.log 2>&1
echo "rerank launches rc=$?"
timeout 200 $T > gpurun_out/${TAG}_small_train.json 2>/dev/null && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1200 --csv --log-file gpurun_out/${TAG}_train_launches.csv $T > gpurun_out/${TAG}_ncu_train.log 2>&1
echo "train launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:gemm_bf16_kernel|t5_self_attn2_kernel" -s 60 -c 6 -f -o gpurun_out/${TAG}_fwd $R > gpurun_out/${TAG}_ncu_full_fwd.log 2>&1
echo "full fwd rc=$?"
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
timeout 100 python /tmp/topk_prof.py && timeout 400 ncu --set full --clock-control none --import-source on -k "regex:topk_warp_kernel" -s 2 -c 1 -f -o gpurun_out/${TAG}_topk python /tmp/topk_prof.py > gpurun_out/${TAG}_ncu_full_topk.log 2>&1
echo "full topk rc=$?"
for f in fwd topk; do ncu -i gpurun_out/${TAG}_$f.ncu-rep --page raw --csv > gpurun_out/${TAG}_${f}_raw.csv 2>/dev/null; done
rm -f gpurun_out/${TAG}_fwd.ncu-rep   # the raw csv is what gets read; keep the small top-k report
ls -la gpurun_out | grep ${TAG}

#!/bin/bash
# top-k v2 with the L2 prefetch (parity + timing), then the attention source-level captures
TAG=${1:-r5b}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_small_kernels.py -q -m gpu -k "topk or empty or argument" > gpurun_out/${TAG}_pytest_topk.log 2>&1; echo "topk pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest_topk.log | cut -c1-300
for v in 2 1 2; do PN_TOPK_WARP=$v timeout 200 python tools/topk_time.py; done 2>&1 | tee gpurun_out/${TAG}_topk_ab.txt
bash tools/gpu_attnprof.sh ${TAG}

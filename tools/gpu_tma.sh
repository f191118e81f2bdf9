#!/bin/bash
# Correctness + same-box A/B of the TMA-fed residual epilogue (PN_GEMM_TMA_RES=0 keeps the register path).
# usage: tools/gpu_tma.sh <tag> [library]
TAG=${1:-tma}; LIB=${2:-pacednegatives_b200/libpn_b200_tma.so}
mkdir -p gpurun_out
export PN_B200_LIB=$PWD/$LIB
timeout 900 python -m pytest tests/test_gpu_t5.py tests/test_gpu_train.py -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/${TAG}_tests.log
for rep in 1 2; do
  for v in 0 1; do
    export PN_GEMM_TMA_RES=$v
    timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_rerank_$v$rep.json 2>/dev/null
    timeout 300 python bench.py --workload train --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_train_$v$rep.json 2>/dev/null
    echo "tma=$v rep$rep rerank $(grep -o '"value": [0-9.]*' gpurun_out/${TAG}_rerank_$v$rep.json | head -1) train $(grep -o '"value": [0-9.]*' gpurun_out/${TAG}_train_$v$rep.json | head -1) hbm-gemm $(grep -o '"hbm_bound_gemms": {[^}]*}' gpurun_out/${TAG}_rerank_$v$rep.json | grep -o '"achieved": [0-9.]*')"
  done
done

#!/bin/bash
# Correctness + same-box A/B of the TMA-fed residual epilogue of the GEMM (PN_GEMM_TMA_RES: 0 off, 1 gated, 2 all).
# usage: tools/gpu_tma.sh <tag>
TAG=${1:-tma}
mkdir -p gpurun_out
for v in 1 2; do   # 1 = default (TMA ring for the epilogue-bound launches), 2 = TMA ring for every residual launch
  PN_GEMM_TMA_RES=$v timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_t5.py tests/test_gpu_train.py -m gpu -x -q > gpurun_out/${TAG}_tests_$v.log 2>&1
  echo "tests (PN_GEMM_TMA_RES=$v) rc=$?"; tail -3 gpurun_out/${TAG}_tests_$v.log
done
for rep in 1 2; do
  for v in 0 1 2; do
    export PN_GEMM_TMA_RES=$v
    timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_rerank_$v$rep.json 2>gpurun_out/${TAG}_rerank_$v$rep.err
    timeout 300 python bench.py --workload train --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_train_$v$rep.json 2>gpurun_out/${TAG}_train_$v$rep.err
    echo "tma=$v rep$rep rerank $(grep -o '"value": [0-9.]*' gpurun_out/${TAG}_rerank_$v$rep.json | head -1) train $(grep -o '"value": [0-9.]*' gpurun_out/${TAG}_train_$v$rep.json | head -1) hbm-gemm $(grep -o '"hbm_bound_gemms": {[^}]*}' gpurun_out/${TAG}_rerank_$v$rep.json | grep -o '"achieved": [0-9.]*')"
  done
done

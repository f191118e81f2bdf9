#!/bin/bash
# fused norm backward: unit tests, training parity, then same-box A/B of the training step
TAG=${1:-r4r}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_attn_bwd.py -x -q -m gpu > gpurun_out/${TAG}_unit.log 2>&1; echo "unit rc=$?"; tail -5 gpurun_out/${TAG}_unit.log | cut -c1-400
timeout 900 python -m pytest tests/test_gpu_train.py -x -q -m gpu > gpurun_out/${TAG}_train.log 2>&1; echo "train rc=$?"; tail -8 gpurun_out/${TAG}_train.log | cut -c1-600
T="python bench.py --workload train --steps 20 --warmup 5 --no-cpu-baseline"
for rep in 1 2; do
  for f in 1 0; do
    timeout 300 $T --set-option fuse_norm_bwd=$f > gpurun_out/${TAG}_train_fuse${f}_${rep}.json 2>gpurun_out/${TAG}_err.log || tail -3 gpurun_out/${TAG}_err.log
    python - <<PY
import json
j=json.load(open("gpurun_out/${TAG}_train_fuse${f}_${rep}.json"))
print("fuse=${f} rep=${rep}", round(j["ms_per_step"],2), "ms", round(j["value"],1), "q/s", j["clocks"]["sm_mhz"])
PY
  done
done

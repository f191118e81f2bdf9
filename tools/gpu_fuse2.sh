#!/bin/bash
# fused norm backward, second pass: tests, A/B (ring depth 3 / 4 / unfused), dropout-regime timing, launch list
TAG=${1:-r4s}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attn_bwd.py tests/test_gpu_gemm.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/${TAG}_unit.log 2>&1; echo "unit rc=$?"; tail -2 gpurun_out/${TAG}_unit.log | cut -c1-300
timeout 900 python -m pytest tests/test_gpu_train.py -x -q -m gpu > gpurun_out/${TAG}_train.log 2>&1; echo "train rc=$?"; tail -2 gpurun_out/${TAG}_train.log | cut -c1-300
T="python bench.py --workload train --steps 20 --warmup 5 --no-cpu-baseline"
run() { # name, env..., extra args
  name=$1; shift
  env "$@" timeout 300 $T $EXTRA > gpurun_out/${TAG}_${name}.json 2>gpurun_out/${TAG}_err.log || tail -3 gpurun_out/${TAG}_err.log
  python - <<PY
import json
j=json.load(open("gpurun_out/${TAG}_${name}.json"))
r=j.get("roofline",{})
print("${name}", round(j["ms_per_step"],2), "ms", round(j["value"],1), "q/s", j["clocks"]["sm_mhz"], "gemm_share", round(r.get("share_of_step",0),3), "attn_bwd", round(r.get("attn_bwd_share",0),3), round(r.get("attn_bwd_tflops",0),1))
PY
}
for rep in 1; do
  EXTRA="--set-option fuse_norm_bwd=1" run fuse_b3_$rep PN_NORM_BUFS=3
  EXTRA="--set-option fuse_norm_bwd=1" run fuse_b4_$rep PN_NORM_BUFS=4
  EXTRA="--set-option fuse_norm_bwd=0" run nofuse_$rep PN_NORM_BUFS=3
done
EXTRA="--train-dropout 0.1" run drop01 PN_NORM_BUFS=3
EXTRA="--train-dropout 0.1 --set-option fuse_norm_bwd=0" run drop01_nofuse PN_NORM_BUFS=3


#!/bin/bash
# full suite + A/B of the norm-backward ring depth + dropout regime + launch lists
TAG=${1:-r4x}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest.log | cut -c1-300
run() { # name, args, env...
  name=$1; shift; wl=$1; shift
  env "$@" timeout 400 python bench.py $wl --no-cpu-baseline --no-library-baseline > gpurun_out/${TAG}_${name}.json 2>gpurun_out/${TAG}_err.log || tail -3 gpurun_out/${TAG}_err.log
  python - <<PY
import json
j=json.load(open("gpurun_out/${TAG}_${name}.json")); r=j.get("roofline",{})
print("${name}", round(j["ms_per_step"],2), "ms", round(j["value"],1), j["unit"], j["clocks"]["sm_mhz"], "gemm", round(r.get("achieved",0)), "attn_bwd", round(r.get("attn_bwd_tflops",0),1), "attn", round(r.get("attn_tflops",0),1))
PY
}
for rep in 1 2; do
  run train_b3_$rep "--workload train --steps 20 --warmup 5" PN_NORM_BUFS=3
  run train_b2_$rep "--workload train --steps 20 --warmup 5" PN_NORM_BUFS=2
done
run drop_b3 "--workload train --steps 10 --warmup 3 --train-dropout 0.1" PN_NORM_BUFS=3
run drop_b2 "--workload train --steps 10 --warmup 3 --train-dropout 0.1" PN_NORM_BUFS=2
run drop_noxattn "--workload train --steps 10 --warmup 3 --train-dropout 0.1 --set-option xattn=0" PN_NORM_BUFS=3
run rerank "--workload rerank --steps 10 --warmup 3" PN_NORM_BUFS=3
T="python bench.py --workload train --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1300 --csv --log-file gpurun_out/${TAG}_train_launches.csv $T > gpurun_out/${TAG}_ncu_train.log 2>&1
echo "launches rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/${TAG}_drop_launches.csv $T --train-dropout 0.1 > gpurun_out/${TAG}_ncu_drop.log 2>&1
echo "drop launches rc=$?"

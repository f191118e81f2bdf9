#!/bin/bash
# programmatic dependent launch of the GEMMs + deeper mask ring: tests, A/B, launch list
TAG=${1:-r4u}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_t5.py -x -q -m gpu > gpurun_out/${TAG}_unit.log 2>&1; echo "unit rc=$?"; tail -2 gpurun_out/${TAG}_unit.log | cut -c1-300
timeout 900 python -m pytest tests/test_gpu_train.py -x -q -m gpu > gpurun_out/${TAG}_train.log 2>&1; echo "train rc=$?"; tail -2 gpurun_out/${TAG}_train.log | cut -c1-300
run() { # name, workload args, env...
  name=$1; shift; wl=$1; shift
  env "$@" timeout 400 python bench.py $wl --no-cpu-baseline --no-library-baseline > gpurun_out/${TAG}_${name}.json 2>gpurun_out/${TAG}_err.log || tail -3 gpurun_out/${TAG}_err.log
  python - <<PY
import json
j=json.load(open("gpurun_out/${TAG}_${name}.json"))
print("${name}", round(j["ms_per_step"],2), "ms", round(j["value"],1), j["unit"], j["clocks"]["sm_mhz"])
PY
}
for rep in 1 2; do
  run train_pdl1_$rep "--workload train --steps 20 --warmup 5" PN_GEMM_PDL=1
  run train_pdl0_$rep "--workload train --steps 20 --warmup 5" PN_GEMM_PDL=0
done
run rerank_pdl1 "--workload rerank --steps 10 --warmup 3" PN_GEMM_PDL=1
run rerank_pdl0 "--workload rerank --steps 10 --warmup 3" PN_GEMM_PDL=0
run rerank_pdl1b "--workload rerank --steps 10 --warmup 3" PN_GEMM_PDL=1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1300 --csv --log-file gpurun_out/${TAG}_train_launches.csv python bench.py --workload train --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_train.log 2>&1
echo "launches rc=$?"

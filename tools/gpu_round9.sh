#!/bin/bash
# residual-epilogue ring depth 5 (4 pipeline stages) against 7 (3 stages): isolated launches, GEMM parity, step A B A B
TAG=${1:-r5d}
mkdir -p gpurun_out
B=$PWD/pacednegatives_b200/libpn_b200_b.so
for v in A B A B; do
  if [ $v = B ]; then export PN_B200_LIB=$B; else unset PN_B200_LIB; fi
  echo "== $v"; timeout 200 python tools/resid_epi.py 2>&1 | cut -c30-200
done | tee gpurun_out/${TAG}_resid_epi_ab.txt
export PN_B200_LIB=$B
timeout 600 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu > gpurun_out/${TAG}_pytest_gemm_b.log 2>&1; echo "gemm pytest (B) rc=$?"; tail -2 gpurun_out/${TAG}_pytest_gemm_b.log | cut -c1-200
unset PN_B200_LIB
for rep in 1 2; do
  for v in A B; do
    if [ $v = B ]; then export PN_B200_LIB=$B; else unset PN_B200_LIB; fi
    python bench.py --workload rerank --steps 5 --warmup 3 --no-cpu-baseline --no-library-baseline > gpurun_out/${TAG}_rerank_$v$rep.json 2>/dev/null
    python bench.py --workload train --steps 10 --warmup 3 --no-cpu-baseline --no-library-baseline > gpurun_out/${TAG}_train_$v$rep.json 2>/dev/null
    python - <<PY
import json
for w in ("rerank", "train"):
    j = json.load(open("gpurun_out/${TAG}_%s_$v$rep.json" % w)); r = j.get("roofline", {})
    print("$v$rep", w, round(j["ms_per_step"], 2), "ms", round(j["value"], 1), j["unit"], j["clocks"]["sm_mhz"], "MHz gemm", round(r.get("achieved", 0)), "hbm-class", round(r.get("hbm_gemm_gbs", 0)), "share", round(r.get("hbm_gemm_share", 0), 3))
PY
  done
done | tee gpurun_out/${TAG}_step_ab.txt

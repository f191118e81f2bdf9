#!/bin/bash
# attention forward with the constant-bias sub-blocks (new) against the bias window from shared memory everywhere (old = build B, -DPN_ATTN_CONST_BIAS=0): parity, isolated timing and timeline (new vs previous build on one box),
# then the re-rank and training steps A B A B
TAG=${1:-r5c}
mkdir -p gpurun_out
OLD=$PWD/pacednegatives_b200/libpn_b200_b.so
timeout 900 python -m pytest tests/test_gpu_t5.py tests/test_gpu_attn_bwd.py tests/test_gpu_train.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/${TAG}_pytest_attn.log 2>&1; echo "attn pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest_attn.log | cut -c1-300
for v in new old new old; do
  if [ $v = old ]; then export PN_B200_LIB=$OLD; else unset PN_B200_LIB; fi
  echo "== $v"; DC=91 timeout 200 python tools/attn_bench.py 2>&1 | tail -3 | cut -c1-200
done | tee gpurun_out/${TAG}_attn_bench_ab.txt
unset PN_B200_LIB
timeout 100 python tools/attn_dbg.py > gpurun_out/${TAG}_attn_timeline.txt 2>&1; echo "timeline rc=$?"
for rep in 1 2; do
  for v in new old; do
    if [ $v = old ]; then export PN_B200_LIB=$OLD; else unset PN_B200_LIB; fi
    python bench.py --workload rerank --steps 5 --warmup 3 --no-cpu-baseline --no-library-baseline > gpurun_out/${TAG}_rerank_$v$rep.json 2>/dev/null
    python bench.py --workload train --steps 10 --warmup 3 --no-cpu-baseline --no-library-baseline > gpurun_out/${TAG}_train_$v$rep.json 2>/dev/null
    python - <<PY
import json
for w in ("rerank", "train"):
    j = json.load(open("gpurun_out/${TAG}_%s_$v$rep.json" % w)); r = j.get("roofline", {})
    print("$v$rep", w, round(j["ms_per_step"], 2), "ms", round(j["value"], 1), j["unit"], j["clocks"]["sm_mhz"], "MHz attn", round(r.get("attn_tflops", 0), 1), "share", round(r.get("attn_share", 0), 3))
PY
  done
done | tee gpurun_out/${TAG}_step_ab.txt

#!/bin/bash
# attention backward (double-buffered resident tiles) + hoisted dropout hash in GEMM epilogues: tests, timing, captures
TAG=${1:-r4w}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attn_bwd.py tests/test_zz_no_alloc.py tests/test_gpu_gemm.py -x -q -m gpu > gpurun_out/${TAG}_unit.log 2>&1; echo "unit rc=$?"; tail -2 gpurun_out/${TAG}_unit.log | cut -c1-300
timeout 900 python -m pytest tests/test_gpu_train.py -x -q -m gpu > gpurun_out/${TAG}_train.log 2>&1; echo "train rc=$?"; tail -2 gpurun_out/${TAG}_train.log | cut -c1-300
T="python bench.py --workload train --steps 1 --warmup 3 --no-cpu-baseline"
for rep in 1 2; do
  for dr in 0.0 0.1; do
  timeout 300 python bench.py --workload train --steps 20 --warmup 5 --no-cpu-baseline --train-dropout $dr > gpurun_out/${TAG}_train_${dr}_$rep.json 2>gpurun_out/${TAG}_err.log || tail -3 gpurun_out/${TAG}_err.log
  python - <<PY
import json
j=json.load(open("gpurun_out/${TAG}_train_${dr}_$rep.json")); r=j["roofline"]
print("train dropout ${dr}", round(j["ms_per_step"],2), "ms", round(j["value"],1), j["clocks"]["sm_mhz"], "attn_bwd", round(r.get("attn_bwd_share",0),3), round(r.get("attn_bwd_tflops",0),1))
PY
  done
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1300 --csv --log-file gpurun_out/${TAG}_train_launches.csv $T > gpurun_out/${TAG}_ncu_train.log 2>&1
echo "launches rc=$?"
cap() { # name, regex, skip
  timeout 500 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$2" -s $3 -c 1 -f -o gpurun_out/${TAG}_$1 $T > gpurun_out/${TAG}_ncu_$1.log 2>&1
  echo "$1 rc=$?"
  ncu -i gpurun_out/${TAG}_$1.ncu-rep --page raw --csv > gpurun_out/${TAG}_$1_raw.csv 2>/dev/null
  ncu -i gpurun_out/${TAG}_$1.ncu-rep --page source --csv > gpurun_out/${TAG}_$1_source.csv 2>/dev/null
  ls -la gpurun_out/${TAG}_$1.ncu-rep | cut -c20-80
}
cap mask 'bool.1, .int.0>' 64
cap normres 'int.3>' 50
cap attnbwd0 't5_attn_bwd_kernel<.int.0>' 25

#!/bin/bash
TAG=${1:-r1w}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
tail -4 gpurun_out/${TAG}_tests.log
python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_rerank.json 2> gpurun_out/${TAG}_bench_rerank.err; echo "rerank rc=$?"
python bench.py --workload select --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_select.json 2> gpurun_out/${TAG}_bench_select.err; echo "select rc=$?"
tail -c 1500 gpurun_out/${TAG}_bench_select.json
SMALL_R="python bench.py --steps 1 --warmup 3 --queries-per-step 1 --no-cpu-baseline"
SMALL_T="python bench.py --workload train --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 600 --csv --log-file gpurun_out/${TAG}_rerank_launches.csv \
  $SMALL_R > gpurun_out/${TAG}_ncu_rerank.log 2>&1
echo "ncu rerank launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:t5_attn_bwd_kernel" -s 24 -c 2 \
  -f -o gpurun_out/${TAG}_attnbwd $SMALL_T > gpurun_out/${TAG}_ncu_attnbwd.log 2>&1
echo "ncu attn bwd rc=$?"
timeout 900 ncu --set full --clock-control none -k "regex:gemm_bf16_kernel|norm_bwd_kernel" -s 434 -c 10 \
  -f -o gpurun_out/${TAG}_encbwd $SMALL_T > gpurun_out/${TAG}_ncu_encbwd.log 2>&1
echo "ncu enc bwd rc=$?"
for f in attnbwd encbwd; do
  ncu -i gpurun_out/${TAG}_$f.ncu-rep --page raw --csv > gpurun_out/${TAG}_${f}_raw.csv 2>/dev/null
done
rm -f gpurun_out/${TAG}_encbwd.ncu-rep
ls -la gpurun_out | grep ${TAG}

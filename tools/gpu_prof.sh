#!/bin/bash
# One gpurun call: new GPU tests, then ncu launch lists + full captures of the headline kernels.
TAG=${1:-r1t}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_wrappers.py tests/test_gpu_train.py tests/test_gpu_t5.py -m gpu -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
tail -15 gpurun_out/${TAG}_tests.log
SMALL_R="python bench.py --steps 1 --warmup 3 --queries-per-step 1 --no-cpu-baseline"
SMALL_T="python bench.py --workload train --steps 1 --warmup 3 --no-cpu-baseline"
$SMALL_T > gpurun_out/${TAG}_small_train.json 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1100 --csv --log-file gpurun_out/${TAG}_train_launches.csv \
  $SMALL_T > gpurun_out/${TAG}_ncu_train.log 2>&1
echo "ncu train launches rc=$?"
$SMALL_R > gpurun_out/${TAG}_small_rerank.json 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 2800 -c 1000 --csv --log-file gpurun_out/${TAG}_rerank_launches.csv \
  $SMALL_R > gpurun_out/${TAG}_ncu_rerank.log 2>&1
echo "ncu rerank launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:gemm_bf16_kernel|t5_self_attn2_kernel" -s 40 -c 6 \
  -f -o gpurun_out/${TAG}_fwd $SMALL_R > gpurun_out/${TAG}_ncu_full_fwd.log 2>&1
echo "ncu full fwd rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:gemm_bf16_kernel|t5_attn_bwd_kernel|norm_bwd_kernel" -s 330 -c 10 \
  -f -o gpurun_out/${TAG}_bwd $SMALL_T > gpurun_out/${TAG}_ncu_full_bwd.log 2>&1
echo "ncu full bwd rc=$?"
for f in fwd bwd; do
  ncu -i gpurun_out/${TAG}_$f.ncu-rep --page raw --csv > gpurun_out/${TAG}_${f}_raw.csv 2>/dev/null
done
ls -la gpurun_out

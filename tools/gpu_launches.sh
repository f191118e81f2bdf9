#!/bin/bash
# ncu launch lists (gpu__time_duration only) of one re-rank step and one training step.  usage: tools/gpu_launches.sh <tag>
TAG=${1:-r3d}
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 600 --csv --log-file gpurun_out/${TAG}_rerank_launches.csv \
  python bench.py --steps 1 --warmup 3 --queries-per-step 1 --no-cpu-baseline > gpurun_out/${TAG}_ncu_rerank.log 2>&1
echo "ncu rerank launches rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1100 --csv --log-file gpurun_out/${TAG}_train_launches.csv \
  python bench.py --workload train --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_train.log 2>&1
echo "ncu train launches rc=$?"

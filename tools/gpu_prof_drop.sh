#!/bin/bash
# launch list of the training step in the dropout regime (p = 0.1, 4 forwards): where the extra time goes
TAG=${1:-r4q}
mkdir -p gpurun_out
T="python bench.py --workload train --train-dropout 0.1 --steps 1 --warmup 3 --no-cpu-baseline"
timeout 300 $T > gpurun_out/${TAG}_drop_train.json 2>gpurun_out/${TAG}_drop_train.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/${TAG}_drop_launches.csv $T > gpurun_out/${TAG}_ncu_drop.log 2>&1
echo "launches rc=$?"
wc -l gpurun_out/${TAG}_drop_launches.csv

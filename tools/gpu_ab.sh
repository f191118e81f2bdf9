#!/bin/bash
# A/B of two builds of the library on ONE box (box-to-box clocks differ by several percent): A B A B.
# usage: tools/gpu_ab.sh <tag> <path to build B>
TAG=${1:-ab}; B=${2:-pacednegatives_b200/libpn_b200_b.so}
mkdir -p gpurun_out
for rep in 1 2; do
  for v in A B; do
    if [ $v = B ]; then export PN_B200_LIB=$PWD/$B; else unset PN_B200_LIB; fi
    python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_rerank_$v$rep.json 2>/dev/null
    python bench.py --workload train --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_train_$v$rep.json 2>/dev/null
    echo "$v$rep rerank $(grep -o '"value": [0-9.]*' gpurun_out/${TAG}_rerank_$v$rep.json | head -1) train $(grep -o '"value": [0-9.]*' gpurun_out/${TAG}_train_$v$rep.json | head -1)"
  done
done

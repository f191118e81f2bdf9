#!/bin/bash
# One gpurun call: GPU tests, a short bench, and the ncu launch list of the same bench command.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r1_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r1_tests.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r1_bench.json
python bench.py --steps 1 --warmup 3 --queries-per-step 1 --no-cpu-baseline > gpurun_out/r1_bench_small.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1_launches.csv \
  python bench.py --steps 1 --warmup 3 --queries-per-step 1 --no-cpu-baseline > gpurun_out/r1_ncu.log 2>&1
echo "ncu rc=$?"

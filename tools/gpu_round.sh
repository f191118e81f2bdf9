#!/bin/bash
# One gpurun call: GPU tests, a short bench, the ncu launch list of a small bench command and one
# ncu --set full capture of the top kernels of that same command.  Usage: tools/gpu_round.sh <tag>
TAG=${1:-r1}
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
tail -3 gpurun_out/${TAG}_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/${TAG}_bench.json
SMALL="python bench.py --steps 1 --warmup 3 --queries-per-step 1 --no-cpu-baseline"
$SMALL > gpurun_out/${TAG}_bench_small.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 2800 -c 1000 --csv --log-file gpurun_out/${TAG}_launches.csv \
  $SMALL > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu launches rc=$?"
$SMALL > gpurun_out/${TAG}_bench_small2.json 2>&1 && \
ncu --set full --clock-control none --import-source on -k "regex:gemm_bf16_kernel|t5_self_attn_kernel" -s 30 -c 5 \
  -f -o gpurun_out/${TAG}_prof $SMALL > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out

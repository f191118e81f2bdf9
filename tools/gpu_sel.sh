#!/bin/bash
TAG=${1:-r1x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_small_kernels.py tests/test_gpu_t5.py -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
tail -12 gpurun_out/${TAG}_tests.log
python bench.py --workload select --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_select.json 2> gpurun_out/${TAG}_bench_select.err; echo "select rc=$?"
tail -c 900 gpurun_out/${TAG}_bench_select.json

#!/bin/bash
TAG=${1:-r1z}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_train.py tests/test_gpu_pairwrapper.py tests/test_gpu_t5.py tests/test_gpu_wrappers.py -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
tail -4 gpurun_out/${TAG}_tests.log
python bench.py --workload train --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_train.json 2> gpurun_out/${TAG}_bench_train.err; echo "train rc=$?"
head -c 400 gpurun_out/${TAG}_bench_train.json

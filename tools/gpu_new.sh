#!/bin/bash
TAG=${1:-r1s}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_wrappers.py tests/test_gpu_train.py tests/test_gpu_t5.py -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
tail -30 gpurun_out/${TAG}_tests.log
timeout 600 python bench.py --workload train --model large --steps 3 --warmup 3 > gpurun_out/${TAG}_bench_train_large.json 2> gpurun_out/${TAG}_bench_train_large.err; echo "large rc=$?"
tail -c 3000 gpurun_out/${TAG}_bench_train_large.json; tail -5 gpurun_out/${TAG}_bench_train_large.err
nvidia-smi --query-gpu=memory.used,memory.total --format=csv

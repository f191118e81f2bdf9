#!/bin/bash
# Quick regression: model-level GPU tests + both headline benches without the CPU baseline.  Usage: tools/gpu_quick.sh <tag>
TAG=${1:-q}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_train.py tests/test_gpu_t5.py tests/test_gpu_pairwrapper.py -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
tail -2 gpurun_out/${TAG}_tests.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_rerank.json 2> gpurun_out/${TAG}_bench_rerank.err; echo "rerank rc=$?"
python bench.py --workload train --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_train.json 2> gpurun_out/${TAG}_bench_train.err; echo "train rc=$?"
for f in rerank train; do grep -o '"value": [0-9.]*, "unit": "[a-z/]*", "n_gpus": 1, "steps": [0-9]*, "warmup": 3, "ms_per_step": [0-9.]*' gpurun_out/${TAG}_bench_$f.json | head -1; done

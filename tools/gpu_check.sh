#!/bin/bash
# One gpurun call: GPU tests, smoke, both headline benches.  Usage: tools/gpu_check.sh <tag>
TAG=${1:-r1}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
tail -3 gpurun_out/${TAG}_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_rerank.json 2> gpurun_out/${TAG}_bench_rerank.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/${TAG}_bench_rerank.json
python bench.py --workload train --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_train.json 2> gpurun_out/${TAG}_bench_train.err; echo "train rc=$?"
tail -c 2500 gpurun_out/${TAG}_bench_train.json

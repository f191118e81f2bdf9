#!/bin/bash
# Round-end check: every GPU test, smoke, and the bench lines of BASELINE configs[1..4].  Usage: tools/gpu_final.sh <tag>
TAG=${1:-r2d}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
tail -3 gpurun_out/${TAG}_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log
python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_rerank.json 2> gpurun_out/${TAG}_bench_rerank.err; echo "rerank rc=$?"
python bench.py --workload train --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_train.json 2> gpurun_out/${TAG}_bench_train.err; echo "train rc=$?"
python bench.py --workload train --model large --steps 3 --warmup 3 > gpurun_out/${TAG}_bench_train_large.json 2> gpurun_out/${TAG}_bench_train_large.err; echo "large rc=$?"
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/${TAG}_bench_rerank_reference.json 2> gpurun_out/${TAG}_bench_rerank_reference.err; echo "ref rc=$?"
python bench.py --workload select --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_select.json 2> gpurun_out/${TAG}_bench_select.err; echo "select rc=$?"
for f in rerank train train_large rerank_reference select; do head -c 330 gpurun_out/${TAG}_bench_$f.json; echo; done

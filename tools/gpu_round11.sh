#!/bin/bash
# (pytest -m gpu -x, smoke, default bench, reference arm) and the ncu launch lists of one re-rank / training step
TAG=${1:-r5f}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
head -c 600 gpurun_out/${TAG}_bench.json; echo
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "reference rc=$?"
head -c 300 gpurun_out/${TAG}_bench_reference.json; echo
R="python bench.py --workload rerank --steps 1 --warmup 3 --queries-per-step 1 --no-cpu-baseline --no-library-baseline"
T="python bench.py --workload train --steps 1 --warmup 3 --no-cpu-baseline --no-library-baseline"
timeout 200 $R > gpurun_out/${TAG}_small_rerank.json 2>/dev/null && \
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 600 --csv --log-file gpurun_out/${TAG}_rerank_launches.csv $R > gpurun_out/${TAG}_ncu_rerank.log 2>&1
echo "rerank launches rc=$?"
timeout 200 $T > gpurun_out/${TAG}_small_train.json 2>/dev/null && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1300 --csv --log-file gpurun_out/${TAG}_train_launches.csv $T > gpurun_out/${TAG}_ncu_train.log 2>&1
echo "train launches rc=$?"
python tools/launch_summary.py gpurun_out/${TAG}_rerank_launches.csv > gpurun_out/${TAG}_rerank_launches_summary.txt 2>&1
python tools/launch_summary.py gpurun_out/${TAG}_train_launches.csv > gpurun_out/${TAG}_train_launches_summary.txt 2>&1
head -8 gpurun_out/${TAG}_rerank_launches_summary.txt

#!/bin/bash
TAG=${1:-r2h}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_small_kernels.py tests/test_gpu_train.py tests/test_gpu_pairwrapper.py tests/test_gpu_wrappers.py -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
tail -3 gpurun_out/${TAG}_tests.log
python bench.py --workload select --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_select.json 2> gpurun_out/${TAG}_bench_select.err; echo "select rc=$?"
python bench.py --workload train --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_train.json 2> gpurun_out/${TAG}_bench_train.err; echo "train rc=$?"
head -c 250 gpurun_out/${TAG}_bench_train.json; echo
python - gpurun_out/${TAG}_bench_select.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
for k in d['kernels']: print(k['kernel'], round(k['ms'],3), round(k['achieved_gbs'],1), round(k['frac_of_hbm_peak'],3))
PY

#!/bin/bash
# usage: tools/gpu_t.sh <tag> <pytest args...>: selected GPU tests, output to gpurun_out/<tag>_pytest.log
TAG=$1; shift
mkdir -p gpurun_out
timeout 1500 python -m pytest "$@" -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
grep -E "^E  |passed|failed|FAILED" gpurun_out/${TAG}_pytest.log | cut -c1-400 | head -60

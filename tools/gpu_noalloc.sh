#!/bin/bash
# the no-alloc entry point's test alone in a fresh process, then after the whole GPU suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/noalloc_suite.log 2>&1; echo "suite rc=$?"; tail -6 gpurun_out/noalloc_suite.log

#!/bin/bash
# Source-level ncu captures (stall samples per SASS instruction) of the encoder self-attention forward and of the two
# backward passes in isolation, the clock-stamp timeline of CTA 0, and the isolated timings.
TAG=${1:-r5b}
mkdir -p gpurun_out
timeout 200 python tools/attn_bench.py > gpurun_out/${TAG}_attn_bench.txt 2>&1; echo "attn_bench rc=$?"; tail -3 gpurun_out/${TAG}_attn_bench.txt | cut -c1-200
timeout 100 python tools/attn_dbg.py > gpurun_out/${TAG}_attn_timeline.txt 2>&1; echo "timeline rc=$?"
cap() { # name, kernel regex, skip, command...
  name=$1; rx=$2; skip=$3; shift 3
  timeout 400 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$rx" -s $skip -c 1 -f -o gpurun_out/${TAG}_$name "$@" > gpurun_out/${TAG}_ncu_$name.log 2>&1
  echo "$name rc=$?"
  ncu -i gpurun_out/${TAG}_$name.ncu-rep --page raw --csv > gpurun_out/${TAG}_${name}_raw.csv 2>/dev/null
  ncu -i gpurun_out/${TAG}_$name.ncu-rep --page source --csv > gpurun_out/${TAG}_${name}_source.csv 2>/dev/null
  ls -la gpurun_out/${TAG}_$name.ncu-rep | cut -c20-90
}
timeout 100 python tools/attn_prof.py && cap attnfwd 't5_self_attn2_kernel' 2 python tools/attn_prof.py
ITERS=2 timeout 100 python tools/attn_bwd_bench.py > gpurun_out/${TAG}_attn_bwd_bench.txt 2>&1 && {
  ITERS=2 cap attnbwd0 't5_attn_bwd_kernel<.int.0>' 1 python tools/attn_bwd_bench.py
  ITERS=2 cap attnbwd1 't5_attn_bwd_kernel<.int.1>' 1 python tools/attn_bwd_bench.py
}
tail -3 gpurun_out/${TAG}_attn_bwd_bench.txt | cut -c1-200
ls -la gpurun_out | grep ${TAG}

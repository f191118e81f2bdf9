#!/bin/bash
# A/B on one box: encoder data gradients to norm_bwd as bf16 (default) or fp32; then the select sweep
TAG=${1:-r4m}
mkdir -p gpurun_out
for rep in 1 2; do for v in 1 0; do
  timeout 300 python bench.py --workload train --steps 10 --warmup 3 --no-cpu-baseline --set-option dgrad_bf16=$v > gpurun_out/${TAG}_train_g$v$rep.json 2>/dev/null
  echo "dgrad_bf16=$v rep=$rep $(grep -o '"value": [0-9.]*' gpurun_out/${TAG}_train_g$v$rep.json | head -1) $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/${TAG}_train_g$v$rep.json | head -1) $(grep -o '"hbm_gemm_gbs": [0-9.]*' gpurun_out/${TAG}_train_g$v$rep.json)"
done; done
timeout 600 python bench.py --workload select --sweep --steps 10 > gpurun_out/${TAG}_select_sweep.json 2> gpurun_out/${TAG}_select_sweep.err; echo "sweep rc=$?"; head -c 600 gpurun_out/${TAG}_select_sweep.json

#!/bin/bash
# ncu --set full of one encoder layer of the re-rank step on the final build (DRAM traffic per launch of the tensor-class
# GEMMs -> profiles/traffic.json); the command first runs without ncu
TAG=${1:-r5m}
mkdir -p gpurun_out
R="python bench.py --workload rerank --steps 1 --warmup 3 --queries-per-step 1 --no-cpu-baseline --no-library-baseline"
timeout 200 $R > gpurun_out/${TAG}_small_rerank.json 2>/dev/null; echo "plain rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:gemm_bf16_kernel|t5_self_attn2_kernel" -s 40 -c 6 -f -o gpurun_out/${TAG}_fwd $R > gpurun_out/${TAG}_ncu_full_fwd.log 2>&1
echo "full fwd rc=$?"
ncu -i gpurun_out/${TAG}_fwd.ncu-rep --page raw --csv > gpurun_out/${TAG}_fwd_raw.csv 2>/dev/null
rm -f gpurun_out/${TAG}_fwd.ncu-rep
ls -la gpurun_out | grep ${TAG}

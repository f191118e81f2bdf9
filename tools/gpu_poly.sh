#!/bin/bash
# A/B/C on one box: PN_ATTN_POLY = 1 (default build), 0, 2: isolated attention forward, then the re-rank step, twice
TAG=${1:-r4h}
mkdir -p gpurun_out
for rep in 1 2; do
  for v in 1 0 2; do
    if [ $v = 1 ]; then unset PN_B200_LIB; else export PN_B200_LIB=$PWD/pacednegatives_b200/libpn_b200_p$v.so; fi
    python tools/attn_bench.py 2>&1 | tail -4 > gpurun_out/${TAG}_attn_p${v}_$rep.txt
    python bench.py --workload rerank --steps 5 --warmup 3 --no-cpu-baseline --no-library-baseline > gpurun_out/${TAG}_rerank_p${v}_$rep.json 2>/dev/null
    echo "poly=$v rep=$rep $(grep 'docs=250' gpurun_out/${TAG}_attn_p${v}_$rep.txt | grep -o 'err=[0-9.e-]* ms=[0-9.]* TFLOP/s=[0-9.]*') rerank $(grep -o '"value": [0-9.]*' gpurun_out/${TAG}_rerank_p${v}_$rep.json | head -1) attn $(grep -o '"attn_tflops": [0-9.]*' gpurun_out/${TAG}_rerank_p${v}_$rep.json)"
    grep "scale=6.0\|scale=3.0" gpurun_out/${TAG}_attn_p${v}_$rep.txt | grep -o "S=[0-9]* H=[0-9]* scale=[0-9.]* err=[0-9.e-]*" | tr '\n' ';'; echo
  done
done

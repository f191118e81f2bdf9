#!/bin/bash
# parity of the GEMM changes, then same-box A/B of PN_GEMM_TMA_MASK and PN_GEMM_MMAJOR (1 = default)
TAG=${1:-r4o}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_train.py tests/test_gpu_t5.py -q -m gpu -x > gpurun_out/${TAG}_pytest.log 2>&1; rc=$?; echo "pytest rc=$rc"; tail -3 gpurun_out/${TAG}_pytest.log | cut -c1-300
if [ $rc != 0 ]; then grep -E "^E  " gpurun_out/${TAG}_pytest.log | head -20 | cut -c1-300; exit 1; fi
for rep in 1 2; do
  for cfg in "1 1" "0 1" "1 0"; do
    set -- $cfg
    export PN_GEMM_TMA_MASK=$1 PN_GEMM_MMAJOR=$2
    timeout 200 python bench.py --workload train --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_train_m$1_j$2_$rep.json 2>/dev/null
    echo "train tma_mask=$1 m_major=$2 rep=$rep $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/${TAG}_train_m$1_j$2_$rep.json | head -1) $(grep -o '"value": [0-9.]*' gpurun_out/${TAG}_train_m$1_j$2_$rep.json | head -1)"
  done
  for j in 1 0; do
    export PN_GEMM_TMA_MASK=1 PN_GEMM_MMAJOR=$j
    timeout 200 python bench.py --workload rerank --steps 5 --warmup 3 --no-cpu-baseline --no-library-baseline > gpurun_out/${TAG}_rerank_j${j}_$rep.json 2>/dev/null
    echo "rerank m_major=$j rep=$rep $(grep -o '"value": [0-9.]*' gpurun_out/${TAG}_rerank_j${j}_$rep.json | head -1) gemm $(grep -o '"achieved": [0-9.]*' gpurun_out/${TAG}_rerank_j${j}_$rep.json | head -1)"
  done
done

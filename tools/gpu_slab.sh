#!/bin/bash
# Row-slab launches of the deep-K residual GEMM (PN_GEMM_SLAB = rows per slab, 0 = one launch): parity tests with a
# small slab so that every such launch in the tests is split, then a same-box A/B of the re-rank step.
TAG=${1:-slab}
mkdir -p gpurun_out
PN_GEMM_SLAB=512 timeout 600 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_t5.py tests/test_gpu_train.py -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1
echo "tests (PN_GEMM_SLAB=512) rc=$?"; tail -2 gpurun_out/${TAG}_tests.log
for rep in 1 2; do
  for v in 0 131072; do
    PN_GEMM_SLAB=$v timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_rerank_${v}_$rep.json 2>/dev/null
    echo "slab=$v rep$rep rerank $(grep -o '"value": [0-9.]*' gpurun_out/${TAG}_rerank_${v}_$rep.json | head -1) gemm $(grep -o '"achieved": [0-9.]*' gpurun_out/${TAG}_rerank_${v}_$rep.json | head -1)"
  done
done

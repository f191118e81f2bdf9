#!/bin/bash
# usage: tools/gpu_scale.sh <tag> <ngpus>   (gpurun --gpus N): the NCCL overlap check (no gloo fallback), then the
# default bench (re-rank + LCE training with its gradient all-reduce) on N GPUs
TAG=${1:-r4}; N=${2:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
PN_REQUIRE_NCCL=1 PYTHONPATH=$PWD timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29731 tests/helpers/overlap_check.py > gpurun_out/${TAG}_overlap_nccl.log 2>&1; echo "overlap rc=$?"; grep "overlap check ok" gpurun_out/${TAG}_overlap_nccl.log
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_bench_${N}gpu.err; echo "bench rc=$?"
head -c 3000 gpurun_out/${TAG}_bench_${N}gpu.json; grep -c "NCCL INFO" gpurun_out/${TAG}_bench_${N}gpu.err; grep -m3 -i "nvls\|Connected all" gpurun_out/${TAG}_bench_${N}gpu.err; tail -5 gpurun_out/${TAG}_bench_${N}gpu.err

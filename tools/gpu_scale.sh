#!/bin/bash
# usage: tools/gpu_scale.sh <tag> <ngpus>
TAG=${1:-r1y}; N=${2:-8}
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --workload train --steps 10 --warmup 3 > gpurun_out/${TAG}_train_${N}gpu.json 2> gpurun_out/${TAG}_train_${N}gpu.err; echo "train rc=$?"
tail -c 600 gpurun_out/${TAG}_train_${N}gpu.json | head -c 600; tail -3 gpurun_out/${TAG}_train_${N}gpu.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${TAG}_rerank_${N}gpu.json 2> gpurun_out/${TAG}_rerank_${N}gpu.err; echo "rerank rc=$?"
head -c 300 gpurun_out/${TAG}_rerank_${N}gpu.json; tail -3 gpurun_out/${TAG}_rerank_${N}gpu.err

#!/bin/bash
TAG=${1:-r2l}; N=${2:-2}
mkdir -p gpurun_out
for c in default 2 4 8; do
  if [ "$c" = default ]; then unset NCCL_MAX_CTAS; else export NCCL_MAX_CTAS=$c; fi
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N bench.py --gpus $N --workload train --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_train_${N}gpu_ctas_$c.json 2> gpurun_out/${TAG}_train_${N}gpu_ctas_$c.err
  echo "NCCL_MAX_CTAS=$c $(head -c 200 gpurun_out/${TAG}_train_${N}gpu_ctas_$c.json | grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' | tr '\n' ' ')"
done

#!/bin/bash
# multi-GPU bench check, launched the way the driver does.  usage: gpurun --gpus N -- 'bash tools/gpu_multi.sh N [tag]'
N=${1:-2}; TAG=${2:-m}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/${TAG}_n$N.json 2> gpurun_out/${TAG}_n$N.err; echo "n$N rc=$?"
cut -c1-1200 gpurun_out/${TAG}_n$N.json; grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/${TAG}_n$N.err | tail -n 6

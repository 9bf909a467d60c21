#!/bin/bash
# multi-GPU bench check, launched the way the driver does.  usage: gpurun --gpus N -- 'bash tools/gpu_multi.sh N [tag]'
N=${1:-2}; TAG=${2:-m}
mkdir -p gpurun_out
timeout 600 python bench.py --gpus 1 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_n1.json 2> gpurun_out/${TAG}_n1.err; echo "n1 rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/${TAG}_n$N.json 2> gpurun_out/${TAG}_n$N.err; echo "n$N rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 \
  bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/${TAG}_ref$N.json 2> gpurun_out/${TAG}_ref$N.err; echo "ref rc=$?"
for f in n1 n$N ref$N; do cut -c1-1500 gpurun_out/${TAG}_$f.json; tail -n 4 gpurun_out/${TAG}_$f.err; done

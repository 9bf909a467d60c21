#!/bin/bash
# e2e jitter hunt: several bench runs with stage traces
mkdir -p gpurun_out
for i in 1 2 3; do
timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --e2e-trace 6 > gpurun_out/e2e_$i.json 2> gpurun_out/e2e_$i.err; echo "rc=$?"
grep "e2e" gpurun_out/e2e_$i.err
done

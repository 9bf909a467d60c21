#!/bin/bash
# A/B of GEMM variants on the bench step (device-resident value + per-class table)
mkdir -p gpurun_out
for impl in 2 3; do
ZV_VERBOSE=1 timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --gemm-impl $impl --profile-out gpurun_out/ab_impl$impl.json > gpurun_out/ab_impl$impl.out 2> gpurun_out/ab_impl$impl.err; echo "impl $impl rc=$?"
grep "zv\]" gpurun_out/ab_impl$impl.err | head -3
python - <<PY
import json
d=json.load(open("gpurun_out/ab_impl$impl.json"))
print(d["ms_per_step"])
for k in d["kernels"][:9]: print(k["name"], k["launches"], k["ms"], k["frac"])
PY
done

#!/bin/bash
# Same-box A/B of two builds of the library: ab/old.so vs ab/new.so, alternating.  usage: tools/gpu_ab_lib.sh [rounds]
mkdir -p gpurun_out
for i in $(seq 1 ${1:-2}); do
  for v in old new; do
    cp ab/$v.so zipvoice_b200/libzipvoice_b200.so
    timeout 600 python bench.py --no-cpu-baseline --no-e2e --profile-out gpurun_out/ab_${v}_$i.json > /dev/null 2> gpurun_out/ab_err.log
    python - <<PY
import json
d = json.load(open("gpurun_out/ab_${v}_$i.json"))
print("$v $i", round(d["ms_per_step"], 1), {k["name"]: round(k["ms"], 1) for k in d["kernels"][:8]})
PY
  done
done
cp ab/new.so zipvoice_b200/libzipvoice_b200.so

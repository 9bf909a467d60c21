#!/bin/bash
# Same-box A/B of several builds of the library (ab/<name>.so): a micro-bench and the config-2 step for each.
# usage: tools/gpu_ab_multi.sh "<micro-bench command>" name1 name2 ...
mkdir -p gpurun_out
micro="$1"; shift
cp zipvoice_b200/libzipvoice_b200.so /tmp/zv_keep.so
for v in "$@"; do
  cp ab/$v.so zipvoice_b200/libzipvoice_b200.so
  echo "== $v"
  [ -n "$micro" ] && (timeout 300 $micro 2>&1 | tee gpurun_out/abm_${v}_micro.txt | grep -v "^mode 4")
  timeout 600 python bench.py --no-cpu-baseline --no-e2e --profile-out gpurun_out/abm_${v}.json > /dev/null 2> gpurun_out/abm_err.log
  python - <<PY
import json
d = json.load(open("gpurun_out/abm_${v}.json"))
print("$v", round(d["ms_per_step"], 1), {k["name"]: round(k["ms"], 1) for k in d["kernels"][:9]})
PY
done
cp /tmp/zv_keep.so zipvoice_b200/libzipvoice_b200.so

#!/bin/bash
# quick GPU pass: kernel unit tests + parity + bench (no ncu).  usage: tools/gpu_quick.sh [tag] [pytest -k expr]
TAG=${1:-q}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -p no:cacheprovider ${2:+-k "$2"} > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" > gpurun_out/${TAG}_rc.txt
timeout 600 python bench.py --no-cpu-baseline --profile-out gpurun_out/${TAG}_kernels.json > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?" >> gpurun_out/${TAG}_rc.txt
cat gpurun_out/${TAG}_rc.txt
tail -n 25 gpurun_out/${TAG}_pytest.log
cut -c1-600 gpurun_out/${TAG}_bench.json
tail -n 5 gpurun_out/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_kernels.json"))
    print(d["ms_per_step"], d["whole_step"])
    for k in d["kernels"]: print(k["name"], k["launches"], k["ms"], k["share"], k["achieved"], k["unit"], k["frac"])
except Exception as e: print("no kernels json", e)
PY

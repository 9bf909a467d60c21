#!/bin/bash
# A/B of kernel variants on the bench step (device-resident value + per-class table).
# usage: tools/gpu_ab.sh "ENV1=.. --flag" "ENV2=.."   (each argument: env assignments and/or bench flags)
mkdir -p gpurun_out
i=0
for v in "$@"; do
i=$((i+1))
envs=""; flags=""
for w in $v; do case "$w" in --*|[0-9]*) flags="$flags $w";; *) envs="$envs $w";; esac; done
env $envs ZV_VERBOSE=1 timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e $flags --profile-out gpurun_out/ab_$i.json > gpurun_out/ab_$i.out 2> gpurun_out/ab_$i.err; echo "variant $i [$v] rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/ab_$i.json"))
print(d["ms_per_step"])
for k in d["kernels"][:10]: print(k["name"], k["launches"], k["ms"], k["frac"])
PY
done

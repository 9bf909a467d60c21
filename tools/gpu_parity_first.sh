#!/bin/bash
# small parity tests under a short timeout first; the quick pass only if they pass
TAG=${1:-p}
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_parity.py -x -q -p no:cacheprovider -k "small or euler" > gpurun_out/${TAG}_par.log 2>&1
rc=$?; echo "parity-first rc=$rc"; tail -n 12 gpurun_out/${TAG}_par.log
if [ $rc -eq 0 ]; then bash tools/gpu_quick.sh $TAG; fi

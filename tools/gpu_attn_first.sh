#!/bin/bash
# attention kernel tests under a short timeout first (new kernel variants can hang); quick pass only if they pass
TAG=${1:-a}
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_gpu_kernels.py -k attention -x -q -p no:cacheprovider > gpurun_out/${TAG}_attn.log 2>&1
rc=$?; echo "attention tests rc=$rc"; tail -n 15 gpurun_out/${TAG}_attn.log
if [ $rc -eq 0 ]; then timeout 100 python tools/attn_bench.py 1219; bash tools/gpu_quick.sh $TAG; fi

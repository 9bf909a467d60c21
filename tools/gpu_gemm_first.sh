#!/bin/bash
# GEMM kernel tests under a short timeout first (new kernel variants can hang); the quick pass only if they pass
TAG=${1:-g}
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_kernels.py -k gemm -x -q -p no:cacheprovider > gpurun_out/${TAG}_gemm.log 2>&1
rc=$?; echo "gemm tests rc=$rc"; tail -n 15 gpurun_out/${TAG}_gemm.log
if [ $rc -eq 0 ]; then bash tools/gpu_quick.sh $TAG; fi

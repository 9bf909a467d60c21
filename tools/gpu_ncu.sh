#!/bin/bash
# ncu evidence only: plain run first, then the launch list, then one full capture.  usage: tools/gpu_ncu.sh [tag]
TAG=${1:-n}
mkdir -p gpurun_out
BCMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 $BCMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s ${NCU_SKIP:-6300} -c ${NCU_COUNT:-420} --csv --log-file gpurun_out/${TAG}_launches.csv $BCMD > gpurun_out/${TAG}_ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${NCU_KERNEL:-gemm_tc_kernel|attn_sa_kernel|attn_na_tc_kernel|dwconv_kernel}" -s ${NCU_FULL_SKIP:-800} -c ${NCU_FULL_COUNT:-30} -o gpurun_out/${TAG}_prof -f $BCMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out/

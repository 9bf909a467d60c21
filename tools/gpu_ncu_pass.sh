#!/bin/bash
# ncu evidence for one build (run under gpurun, one GPU): launch list of one NFE of the bench command and a full capture
# of the top kernels, both only after the same command exited 0 without ncu.  Keeps what comes back small (gpurun
# returns at most 64 MiB): the raw pages are exported to CSV on the box and the .ncu-rep is dropped when it is large.
# usage: tools/gpu_ncu_pass.sh [tag]
TAG=${1:-r2}
mkdir -p gpurun_out
BCMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 $BCMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s ${NCU_SKIP:-6300} -c ${NCU_COUNT:-420} --csv \
  --log-file gpurun_out/${TAG}_launches.csv $BCMD > gpurun_out/${TAG}_ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on \
  -k regex:"${NCU_KERNEL:-gemm_tc_kernel|attn_sa2_kernel|attn_na_tc_kernel|dwconv_kernel}" -s ${NCU_FULL_SKIP:-800} \
  -c ${NCU_FULL_COUNT:-16} -o gpurun_out/${TAG}_prof -f $BCMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i gpurun_out/${TAG}_prof.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ls -la gpurun_out/${TAG}_*
SZ=$(stat -c %s gpurun_out/${TAG}_prof.ncu-rep 2>/dev/null || echo 0)
if [ "$SZ" -gt 40000000 ]; then rm -f gpurun_out/${TAG}_prof.ncu-rep; echo "dropped the .ncu-rep ($SZ bytes), kept the CSV"; fi

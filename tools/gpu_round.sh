#!/bin/bash
# One GPU validation pass: parity tests, smoke, bench (N=1), ncu launch list + full capture of the top kernels.
# usage: tools/gpu_round.sh [tag]      (outputs under gpurun_out/<tag>_*)
TAG=${1:-r1}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" > gpurun_out/${TAG}_rc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${TAG}_rc.txt
timeout 600 python bench.py --profile-out gpurun_out/${TAG}_kernels.json > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?" >> gpurun_out/${TAG}_rc.txt
if [ "${NCU:-1}" = "1" ]; then
BCMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 $BCMD > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s ${NCU_SKIP:-6300} -c ${NCU_COUNT:-420} --csv --log-file gpurun_out/${TAG}_launches.csv $BCMD > gpurun_out/${TAG}_ncu_list.log 2>&1
echo "ncu list rc=$?" >> gpurun_out/${TAG}_rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${NCU_KERNEL:-gemm_tc_kernel|attn_sa_kernel|attn_na_tc_kernel|dwconv_kernel}" -s ${NCU_FULL_SKIP:-800} -c ${NCU_FULL_COUNT:-40} -o gpurun_out/${TAG}_prof -f $BCMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?" >> gpurun_out/${TAG}_rc.txt
fi
cat gpurun_out/${TAG}_rc.txt
tail -3 gpurun_out/${TAG}_pytest.log gpurun_out/${TAG}_smoke.log
cat gpurun_out/${TAG}_bench.json | cut -c1-1500
tail -5 gpurun_out/${TAG}_bench.err

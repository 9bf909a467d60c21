#!/bin/bash
# first GPU validation pass: kernel unit parity, then model parity
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout 400 python -m pytest tests/test_gpu_kernels.py -q -k "simt or pos_table or attention" -p no:cacheprovider > gpurun_out/t1_simt_attn.log 2>&1; echo "t1 rc=$?" >> gpurun_out/rc.txt
timeout 300 python -m pytest tests/test_gpu_kernels.py -q -k "tcgen05" -p no:cacheprovider > gpurun_out/t2_tcgen05.log 2>&1; echo "t2 rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m pytest tests/test_gpu_parity.py -q -p no:cacheprovider > gpurun_out/t3_parity.log 2>&1; echo "t3 rc=$?" >> gpurun_out/rc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/t4_smoke.log 2>&1; echo "t4 rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
tail -5 gpurun_out/t1_simt_attn.log gpurun_out/t2_tcgen05.log gpurun_out/t3_parity.log gpurun_out/t4_smoke.log

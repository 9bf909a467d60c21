"""Kernel-only timings of the residual GEMM on the shapes of the B=64 step: fp32 stream (EPI_RESIDUAL, the vocoder's
epilogue) against the s24 stream planes (EPI_RES24, the Zipformer layers'), with both rooflines per shape.
usage: python tools/res24_bench.py [impl ...]      env: ZV_GEMM_STAGES=n caps the ring depth"""
import ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
pk = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))) \
    if os.path.exists("MEASURED_PEAKS.json") else {}
HBM = float(pk.get("hbm_gbs", 6464.3))
TF = float(pk.get("bf16_tflops_sustained", 1438.3))
D = 512
def run(mode, M, K, impl, iters=10):
    ms = C.c_float()
    rc = lib.zv_debug_gemm_bench(mode, impl, M, D, K, 1219, 0, 0, iters, C.byref(ms))
    by = M * (2.0 * K + (6.0 if mode == 8 else 10.0) * D)
    us = ms.value * 1e3
    gbs = by / (ms.value * 1e-3) / 1e9 if ms.value > 0 else 0.0
    tfs = 2.0 * M * D * K / (ms.value * 1e-3) / 1e12 if ms.value > 0 else 0.0
    bound_us = max(by / (HBM * 1e9), 2.0 * M * D * K / (TF * 1e12)) * 1e6
    print(f"mode {mode} impl {impl} M={M} K={K:4d} rc={rc} {us:8.1f} us  {gbs:7.1f} GB/s ({gbs / HBM:5.3f})  "
          f"{tfs:7.1f} TF/s ({tfs / TF:5.3f})  roofline {bound_us:7.1f} us -> {bound_us / us if us else 0:5.3f}")
impls = [int(a) for a in sys.argv[1:]] or [0]
for M in (128 * 1219, 128 * 610):
    for K in (1152, 1536, 1920, 384, 576):
        for impl in impls:
            for mode in (4, 8):
                run(mode, M, K, impl)

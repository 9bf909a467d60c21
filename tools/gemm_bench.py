"""Kernel-only timings of the tcgen05 GEMM on the shapes of the B=64 step (synthetic data) for profiling experiments.
usage: python tools/gemm_bench.py [impl]      env: ZV_GEMM_STAGES=n caps the ring depth"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
BIAS, SWOOSHL, GLU, NONLIN, RESIDUAL, F32 = 0, 1, 2, 3, 4, 5
PEAK = 1414.0
def run(tag, mode, M, N, K, impl=0, bn=0, gate=0, iters=5):
    ms = C.c_float()
    rc = lib.zv_debug_gemm_bench(mode, impl, M, N, K, 1219, gate, bn, iters, C.byref(ms))
    tf = 2.0 * M * N * K / (ms.value * 1e-3) / 1e12 if ms.value > 0 else 0
    print(f"{tag:30s} impl={impl} bn={bn:3d} M={M} N={N:4d} K={K:4d} rc={rc} {ms.value*1e3:8.1f} us {tf:7.1f} TF/s {tf/PEAK:6.3f}")
impl = int(sys.argv[1]) if len(sys.argv) > 1 else 0
M = 128 * 1219
for N in (1152, 1536, 1920):
    run("FF in  bias-only epilogue", BIAS, M, N, 512, impl)
    run("FF in  swooshl", SWOOSHL, M, N, 512, impl)
    run("FF in  swooshl bn=128", SWOOSHL, M, N, 512, impl, bn=128)
run("conv in glu", GLU, M, 1024, 512, impl)
run("nonlin in", NONLIN, M, 1152, 512, impl, gate=768)
run("attn in bias N=272", BIAS, M, 272, 512, impl)
for K in (1152, 1536, 1920, 512, 384, 48):
    run("out residual", RESIDUAL, M, 512, K, impl)

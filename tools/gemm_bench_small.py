"""Residual GEMMs with small K (conv out_proj K=576, NonlinAttention out_proj K=384): impl x block_n sweep."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
RESIDUAL = 4
def run(M, N, K, impl, bn, iters=8):
    ms = C.c_float()
    rc = lib.zv_debug_gemm_bench(RESIDUAL, impl, M, N, K, 1219, 0, bn, iters, C.byref(ms))
    tf = 2.0 * M * N * K / (ms.value * 1e-3) / 1e12 if ms.value > 0 else 0
    gb = M * (K + 2 * N) * 2 / (ms.value * 1e-3) / 1e9 if ms.value > 0 else 0
    print(f"M={M} N={N} K={K:4d} impl={impl} bn={bn:3d} rc={rc} {ms.value*1e3:8.1f} us {tf:7.1f} TF/s {gb:7.0f} GB/s")
for M in (128 * 1219, 128 * 610):
    for K in (576, 384, 1152):
        for impl in (2, 3):
            for bn in (0, 128, 256):
                run(M, 512, K, impl, bn)

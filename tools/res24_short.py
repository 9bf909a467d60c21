"""Kernel-only time of the two short-K residual GEMM shapes (EPI_RES24), for measurement builds."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
for K in (384, 576, 1152, 1536):
    for mode, tag in ((8, "two planes"),):
        ms = C.c_float()
        rc = lib.zv_debug_gemm_bench(mode, 0, 128 * 1219, 512, K, 1219, 0, 0, 10, C.byref(ms))
        print(f"RES24 {tag:17s} M=156032 K={K:4d} rc={rc} {ms.value * 1e3:8.1f} us")

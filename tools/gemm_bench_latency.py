"""Small-M GEMMs of a single utterance (M = 750 and 1500 rows): device time per launch, back-to-back launches.
usage: [ZV_GEMM_TAILSPLIT=1|2] python tools/gemm_bench_latency.py"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
MODES = {"bias": 0, "swooshl": 1, "glu": 2, "nonlin": 3, "residual": 8}   # 8: the s24 residual epilogue of the layers
def run(name, M, N, K, iters=50):
    ms = C.c_float()
    rc = lib.zv_debug_gemm_bench(MODES[name], 0, M, N, K, M, 0, 0, iters, C.byref(ms))
    print(f"{name:9s} M={M:5d} N={N:4d} K={K:4d} rc={rc} {ms.value*1e3:7.2f} us")
def run_nl(M, iters=50):
    ms = C.c_float()
    rc = lib.zv_debug_gemm_bench(3, 0, M, 1152, 512, M, 768, 0, iters, C.byref(ms))
    print(f"nonlin    M={M:5d} N=1152 K= 512 rc={rc} {ms.value*1e3:7.2f} us")
for M in (750, 1500, 3000):
    run("residual", M, 512, 1920)
    run("residual", M, 512, 1536)
    run("residual", M, 512, 1152)
    run("residual", M, 512, 576)
    run("residual", M, 512, 384)
    run("swooshl", M, 1536, 512)
    run("glu", M, 1024, 576)
    run_nl(M)
    run("bias", M, 272, 512)

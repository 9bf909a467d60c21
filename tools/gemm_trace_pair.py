"""ZV_TRACE build: producer / MMA thread cycle split of the FF in_proj + SwooshL GEMM, single-CTA tiles (impl 2) against
CTA pairs (impl 3).  usage: ZV_NVCC_DEFS=-DZV_TRACE python -m zipvoice_b200.build --force; python tools/gemm_trace_pair.py"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
for mode, name in ((1, "swooshl"), (0, "bias")):
    for impl in (2, 3):
        ms = C.c_float()
        rc = lib.zv_debug_gemm_bench(mode, impl, 128 * 1219, 1536, 512, 1219, 0, 0, 1, C.byref(ms))
        torch.cuda.synchronize()
        print(f"== {name} impl {impl} rc={rc} {ms.value * 1e3:8.1f} us", flush=True)

"""Short-K residual GEMM (EPI_RES24) against tile width; run with ZV_GEMM_STAGES=n to cap the operand ring depth."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
for K in (384, 576):
    for bn in (256, 128):
        ms = C.c_float()
        rc = lib.zv_debug_gemm_bench(8, 2, 128 * 1219, 512, K, 1219, 0, bn, 10, C.byref(ms))
        print(f"stages<={os.environ.get('ZV_GEMM_STAGES', 'max')} RES24 M=156032 K={K:4d} bn={bn} rc={rc} {ms.value * 1e3:8.1f} us")

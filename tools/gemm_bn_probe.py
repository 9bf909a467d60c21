import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
for M in (750,):
  for K in (1536, 576):
    for bn in (256, 128, 64, 32):
        ms = C.c_float()
        rc = lib.zv_debug_gemm_bench(4, 2, M, 512, K, M, 0, bn, 50, C.byref(ms))
        print(f"residual M={M} K={K} bn={bn} rc={rc} {ms.value*1e3:7.2f} us")
  for bn in (256, 128, 64, 32):
        ms = C.c_float()
        rc = lib.zv_debug_gemm_bench(0, 2, M, 512, 512, M, 0, bn, 50, C.byref(ms))
        print(f"bias M={M} K=512 N=512 bn={bn} rc={rc} {ms.value*1e3:7.2f} us")

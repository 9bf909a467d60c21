"""ZV_TRACE build only (ZV_NVCC_DEFS=-DZV_TRACE): cycle split of the producer and MMA threads of CTA 0."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
for (M, N, K) in ((750, 512, 1536), (750, 512, 576), (156032, 512, 1536)):
    ms = C.c_float()
    print(f"--- residual M={M} N={N} K={K}", flush=True)
    rc = lib.zv_debug_gemm_bench(4, 2, M, N, K, M if M < 2000 else 1219, 0, 0, 2, C.byref(ms))
    torch.cuda.synchronize()
    print(f"rc={rc} {ms.value*1e3:.2f} us", flush=True)

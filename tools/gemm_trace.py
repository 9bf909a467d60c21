"""ZV_TRACE build only (ZV_NVCC_DEFS=-DZV_TRACE): cycle split of the producer and MMA threads of CTA 0."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
CASES = ((4, 750, 512, 1536, 0), (4, 750, 512, 576, 0), (4, 156032, 512, 1536, 0), (1, 156032, 1536, 512, 0),
         (2, 156032, 1024, 576, 0), (3, 156032, 1152, 512, 768), (4, 156032, 512, 576, 0))
for (mode, M, N, K, gate) in CASES:
    ms = C.c_float()
    print(f"--- mode {mode} M={M} N={N} K={K}", flush=True)
    rc = lib.zv_debug_gemm_bench(mode, 2, M, N, K, M if M < 2000 else 1219, gate, 0, 1, C.byref(ms))
    torch.cuda.synchronize()
    print(f"rc={rc} {ms.value*1e3:.2f} us", flush=True)

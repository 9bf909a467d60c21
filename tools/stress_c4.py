"""Isolated launches of the kernels of the dialog configuration (B' = 2, T = 6563 and its downsampled lengths), many times
each, to find which one fails: prints rc per kernel and shape."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
ms = C.c_float()
def chk(tag, rc):
    err = torch.cuda.synchronize() if False else None
    print(f"{tag:50s} rc={rc} {ms.value * 1e3:9.1f} us", flush=True)
    if rc:
        sys.exit(1)
for L in (6563, 3282, 1641):
    for rep in range(3):
        chk(f"NA nb=2 L={L}", lib.zv_debug_attention_bench(2, L, 4, 1, 384, 0, 0, 5, C.byref(ms)))
        chk(f"SA nb=2 L={L}", lib.zv_debug_attention_bench(2, L, 4, 4, 12, 0, 0, 5, C.byref(ms)))
        chk(f"SA nb=2 L={L} mask", lib.zv_debug_attention_bench(2, L, 4, 4, 12, 1, 0, 5, C.byref(ms)))
    M = 2 * L
    for mode, N, K, gate in ((1, 1424, 512, 1152), (1, 1536, 512, 0), (1, 1920, 512, 0), (2, 1024, 576, 0), (3, 1152, 512, 768),
                             (0, 64, 512, 0), (8, 512, 1152, 0), (8, 512, 1536, 0), (8, 512, 1920, 0), (8, 512, 384, 0),
                             (8, 512, 576, 0)):
        for rep in range(3):
            chk(f"GEMM mode {mode} M={M} N={N} K={K}", lib.zv_debug_gemm_bench(mode, 0, M, N, K, L, gate, 0, 5, C.byref(ms)))
print("all ok")

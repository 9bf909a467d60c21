"""Kernel-only time of the NonlinAttention kernel on the shapes of the B=64 step (measurement builds, no parity)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
for L in (305, 610, 1219):
    ms = C.c_float()
    rc = lib.zv_debug_attention_bench(128, L, 4, 1, 384, 0, 0, 10, C.byref(ms))
    print(f"NA nb=128 L={L:5d} rc={rc} {ms.value:8.4f} ms")

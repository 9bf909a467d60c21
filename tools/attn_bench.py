"""Kernel-only timings of the attention kernels (synthetic data) for profiling experiments."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import _lib
lib = _lib.load()
torch.zeros(1, device="cuda")
def run(tag, nb, L, heads, dv, impl=0, dbg=0, mask=0, iters=5):
    lib.zv_debug_set_attn_impl(impl)
    ms = C.c_float()
    rc = lib.zv_debug_attention_bench(nb, L, 4, heads, dv, mask, dbg, iters, C.byref(ms))
    pairs = nb * L * L
    print(f"{tag:34s} nb={nb} L={L:5d} rc={rc} {ms.value:8.3f} ms  {pairs / ms.value / 1e6:8.1f} G pairs/s")
only = int(sys.argv[1]) if len(sys.argv) > 1 else 0
if only:
    run("NA tcgen05", 128, only, 1, 384, iters=2)
    run("SA", 128, only, 4, 12, iters=2)
    sys.exit(0)
for L in (305, 610, 1219):
    run("NA tcgen05", 128, L, 1, 384)
    run("NA tcgen05 no-V-load", 128, L, 1, 384, dbg=1)
    run("NA tcgen05 no-MMA", 128, L, 1, 384, dbg=2)
    run("NA tcgen05 no-V no-MMA", 128, L, 1, 384, dbg=3)
    run("NA mma.sync", 128, L, 1, 384, impl=1)
    run("SA", 128, L, 4, 12)
    run("SA masked-tail", 128, L, 4, 12, mask=1)
run("NA tcgen05 dialog", 2, 6563, 1, 384)
run("SA dialog", 2, 6563, 4, 12)
lib.zv_debug_set_attn_impl(0)

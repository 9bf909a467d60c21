"""A/B of the two SelfAttention kernels (ZV_SA_IMPL=1: position term on CUDA cores, 2: on tcgen05): accuracy against
the fp32 torch reference of tests/test_gpu_kernels.py and kernel-only time.  Run once per setting of ZV_SA_IMPL."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests.test_gpu_kernels import _lib, _p, ref_attention, rel
lib = _lib()
H, nb, dv = 4, 3, 12
for L in (64, 130, 611, 1219):
    for escale in (1.0, 4.0):
        g = torch.Generator(device="cuda").manual_seed(L)
        rn = lambda *s: torch.randn(*s, device="cuda", generator=g)
        qkp = (rn(nb, L, 288) * 0.7).bfloat16().float()
        E = (rn(H, 2 * L - 1, 4) * escale).contiguous()
        lens = torch.tensor([L, max(1, L // 2), max(1, L - 3)], device="cuda")
        mask = (torch.arange(L, device="cuda")[None] >= lens[:, None]).to(torch.uint8).contiguous()
        V = rn(nb, L, H * dv).bfloat16().float()
        out = torch.zeros(nb, L, H * dv, device="cuda")
        rc = lib.zv_debug_attention(nb, L, H, H, dv, _p(qkp), 288, _p(E), _p(mask), _p(V), H * dv, None, _p(out), None)
        want = ref_attention(qkp, E, mask, V, None, H, H, dv)
        print(f"impl {os.environ.get('ZV_SA_IMPL', '2')} L={L:5d} E x{escale}: rc={rc} rel-L2 {rel(out, want):.3e} "
              f"max-abs {float((out - want).abs().max()):.3e}")
for L in (305, 610, 1219):
    for mask in (0, 1):
        ms = C.c_float()
        rc = lib.zv_debug_attention_bench(128, L, 4, 4, 12, mask, 0, 10, C.byref(ms))
        print(f"impl {os.environ.get('ZV_SA_IMPL', '2')} SA nb=128 L={L:5d} mask={mask} rc={rc} {ms.value:8.4f} ms")
ms = C.c_float()
rc = lib.zv_debug_attention_bench(2, 6563, 4, 4, 12, 0, 0, 5, C.byref(ms))
print(f"impl {os.environ.get('ZV_SA_IMPL', '2')} SA nb=2 L=6563 rc={rc} {ms.value:8.4f} ms")

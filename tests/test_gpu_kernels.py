"""Unit parity of single CUDA kernels through the C ABI debug hooks (GPU only).
Reference = plain torch fp32 on the bf16-rounded operands (these are floating-point kernels)."""
import ctypes as C
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

EPI_BIAS, EPI_SWOOSHL, EPI_GLU, EPI_NONLIN, EPI_RESIDUAL, EPI_F32 = range(6)
EPI_RES24 = 8  # debug-hook number of the residual epilogue on the s24 stream (hi bf16 plane + lo byte plane)


def _lib():
    from zipvoice_b200 import _lib
    return _lib.load()


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def swoosh_l(x):
    return torch.logaddexp(torch.zeros_like(x), x - 4.0) - 0.08 * x - 0.035


def run_gemm(mode, impl, A, W, bias, L=1, gate_cols=0, rowbias=None, res=None, resbias=None, orig=None, bscale=None,
             mask=None):
    M, K = A.shape
    N = W.shape[0]
    split = mode == EPI_SWOOSHL and gate_cols > 0   # FF1 in_proj + attention q | k | p in one launch
    n_out = N // 2 if mode == EPI_GLU else (gate_cols // 2 if mode == EPI_NONLIN else (gate_cols if split else N))
    out = torch.zeros(M, n_out, device="cuda")
    out2 = torch.zeros(M, max(N - gate_cols, 1) if (mode == EPI_NONLIN or split) else N, device="cuda")
    rc = _lib().zv_debug_gemm(mode, impl, M, N, K, L, gate_cols, _p(A), _p(W), _p(bias), _p(rowbias), _p(res),
                              _p(resbias), _p(orig), _p(bscale), _p(mask), _p(out), _p(out2), None)
    assert rc == 0, f"zv_debug_gemm rc={rc}"
    return out, out2


def ref_gemm(mode, A, W, bias, L=1, gate_cols=0, rowbias=None, res=None, resbias=None, orig=None, bscale=None,
             mask=None):
    M = A.shape[0]
    acc = A.bfloat16().float() @ W.bfloat16().float().T + bias
    b = torch.arange(M, device="cuda") // L
    if rowbias is not None and mode not in (EPI_RESIDUAL, EPI_RES24):
        acc = acc + rowbias[b]
    if mode == EPI_BIAS or mode == EPI_F32:
        return acc, None
    if mode == EPI_SWOOSHL and gate_cols > 0:
        # columns >= gate_cols: plain bias, no row bias, stored as fp16
        tail = A.bfloat16().float() @ W.bfloat16().float().T[:, gate_cols:] + bias[gate_cols:]
        return swoosh_l(acc[:, :gate_cols]), tail
    if mode == EPI_SWOOSHL:
        return swoosh_l(acc), None
    if mode == EPI_GLU:
        g = acc.reshape(M, -1, 2, 16)
        o = (g[:, :, 0] * torch.sigmoid(g[:, :, 1])).reshape(M, -1)
        if mask is not None:
            o = o * (1 - mask.reshape(-1).float())[:, None]
        return o, None
    if mode == EPI_NONLIN:
        g = acc[:, :gate_cols].reshape(M, -1, 2, 16)
        return (g[:, :, 1] * torch.tanh(g[:, :, 0])).reshape(M, -1), acc[:, gate_cols:]
    if mode in (EPI_RESIDUAL, EPI_RES24):
        v = acc + res if res is not None else acc
        if resbias is not None:
            v = v + resbias[b]
        if orig is not None:
            v = orig + (v - orig) * bscale
        return v, v
    raise AssertionError


def rel(a, b):
    return float((a - b).norm() / b.norm().clamp_min(1e-20))


CASES = [
    # mode, M, N, K, extra
    (EPI_F32, 300, 512, 320, {}),
    (EPI_F32, 257, 100, 512, {}),
    (EPI_BIAS, 1000, 272, 512, {}),
    (EPI_BIAS, 130, 48, 512, {}),
    (EPI_SWOOSHL, 777, 1152, 512, {"rowbias": True, "L": 111}),
    (EPI_SWOOSHL, 640, 1920, 512, {}),
    (EPI_SWOOSHL, 700, 1424, 512, {"rowbias": True, "L": 100, "gate_cols": 1152}),
    (EPI_SWOOSHL, 40000, 1424, 512, {"rowbias": True, "L": 1219, "gate_cols": 1152}),
    (EPI_SWOOSHL, 333, 656, 192, {"gate_cols": 384}),
    (EPI_GLU, 500, 1024, 512, {"mask": True, "L": 125}),
    (EPI_GLU, 200, 384, 192, {}),
    (EPI_NONLIN, 400, 1152, 512, {"gate_cols": 768}),
    (EPI_NONLIN, 150, 432, 192, {"gate_cols": 288}),
    (EPI_RESIDUAL, 900, 512, 1536, {"resbias": True, "orig": True, "L": 300}),
    (EPI_RESIDUAL, 333, 512, 48, {}),
    (EPI_RESIDUAL, 129, 192, 640, {}),
    (EPI_RESIDUAL, 20000, 512, 384, {"resbias": True, "L": 1219}),
    # several persistent rounds per CTA / CTA pair, odd tile counts, ragged last tile
    (EPI_SWOOSHL, 40100, 1536, 512, {"rowbias": True, "L": 1219}),
    (EPI_GLU, 38000, 1024, 512, {"mask": True, "L": 1000}),
    (EPI_NONLIN, 38500, 1152, 512, {"gate_cols": 768}),
    (EPI_BIAS, 39000, 272, 512, {}),
    (EPI_RESIDUAL, 38300, 512, 1920, {"orig": True}),
    # partial last wave cut into column slices (TileSched in zv_gemm.cu): 150 tiles -> 2 leftover tiles in 2 slices;
    # the quarter-rate residual shape of the bench, 610 tiles -> 18 leftover tiles in 4 slices
    (EPI_SWOOSHL, 3823, 1152, 512, {"rowbias": True, "L": 1000}),
    (EPI_RESIDUAL, 39040, 512, 576, {"resbias": True, "L": 305}),
    # the same residual epilogue on the s24 stream planes (the Zipformer layers); "nores": stack in_proj, no input
    (EPI_RES24, 900, 512, 1536, {"resbias": True, "orig": True, "L": 300}),
    (EPI_RES24, 333, 512, 48, {}),
    (EPI_RES24, 129, 192, 640, {}),
    (EPI_RES24, 300, 512, 320, {"nores": True}),
    (EPI_RES24, 20000, 512, 384, {"resbias": True, "L": 1219}),
    (EPI_RES24, 38300, 512, 1920, {"orig": True}),
    (EPI_RES24, 39040, 512, 576, {"resbias": True, "L": 305}),
]


@pytest.mark.parametrize("impl", [1, 2, 3, 0], ids=["simt", "tcgen05", "tcgen05_pair", "auto"])
@pytest.mark.parametrize("case", CASES, ids=lambda c: f"m{c[0]}_{c[1]}x{c[2]}x{c[3]}")
def test_gemm(case, impl):
    mode, M, N, K, ex = case
    if impl == 1 and M * N * K > 3e9:
        pytest.skip("SIMT checker kept to small shapes")
    g = torch.Generator(device="cuda").manual_seed(mode * 1000 + M)
    rn = lambda *s: torch.randn(*s, device="cuda", generator=g)
    A, W, bias = rn(M, K), rn(N, K) / math.sqrt(K), rn(N) * 0.5
    L = ex.get("L", M)
    nb = (M + L - 1) // L
    kw = dict(L=L, gate_cols=ex.get("gate_cols", 0))
    if ex.get("rowbias"):
        kw["rowbias"] = rn(nb, N)
    if mode in (EPI_RESIDUAL, EPI_RES24):
        kw["res"] = None if ex.get("nores") else rn(M, N)
        if ex.get("resbias"):
            kw["resbias"] = rn(nb, N)
        if ex.get("orig"):
            kw["orig"], kw["bscale"] = rn(M, N), torch.rand(N, device="cuda", generator=g)
    if ex.get("mask"):
        kw["mask"] = (torch.rand(nb * L, device="cuda", generator=g) < 0.3).to(torch.uint8)[: nb * L].contiguous()
    out, out2 = run_gemm(mode, impl, A, W, bias, **kw)
    want, want2 = ref_gemm(mode, A, W, bias, **kw)
    # bf16 outputs: 2^-9 rounding; s24: 16 significant bits (2^-17 per value, on the residual input and the output)
    tol = 2e-5 if mode in (EPI_F32, EPI_RESIDUAL, EPI_RES24) else 6e-3
    assert rel(out, want) < tol, f"rel-L2 {rel(out, want):.3e} max-abs {float((out - want).abs().max()):.3e}"
    if mode == EPI_NONLIN or (mode == EPI_SWOOSHL and ex.get("gate_cols")):
        assert rel(out2, want2) < 6e-3, f"out2 rel-L2 {rel(out2, want2):.3e}"
    if mode in (EPI_RESIDUAL, EPI_RES24):
        assert rel(out2, want2) < 6e-3, f"bf16 shadow rel-L2 {rel(out2, want2):.3e}"


def test_pos_table():
    from oracle import zipformer_oracle as zo
    for L in (5, 64, 305, 1219):
        W = torch.randn(16, 48, generator=torch.Generator().manual_seed(L)) * 0.3
        want = (zo.rel_pos_encoding(L, 48) @ W.T).reshape(2 * L - 1, 4, 4).permute(1, 0, 2).contiguous()
        E = torch.zeros(4, 2 * L - 1, 4, device="cuda")
        rc = _lib().zv_debug_pos_table(_p(W.cuda()), 4, 4, 48, L, _p(E), None)
        assert rc == 0
        err = float((E.cpu() - want).abs().max())
        assert err < 2e-5, f"L={L} max-abs {err:.3e}"


def ref_attention(qkp, E, mask, V, gate, H, heads, dv):
    nb, L, _ = qkp.shape
    q = qkp[..., : H * 32].reshape(nb, L, H, 32).permute(2, 0, 1, 3)
    k = qkp[..., H * 32: 2 * H * 32].reshape(nb, L, H, 32).permute(2, 0, 1, 3)
    p = qkp[..., 2 * H * 32: 2 * H * 32 + 4 * H].reshape(nb, L, H, 4).permute(2, 0, 1, 3)
    s = q @ k.transpose(-1, -2)
    ps = p @ E.permute(0, 2, 1)[:, None]  # (H, nb, L, 2L-1)
    idx = (L - 1) - torch.arange(L, device="cuda")[:, None] + torch.arange(L, device="cuda")[None, :]
    s = s + torch.gather(ps, 3, idx[None, None].expand(H, nb, L, L))
    if mask is not None:
        s = s.masked_fill(mask.bool()[None, :, None, :], -1000.0)
    a = s.softmax(-1)[:heads]
    v = V[..., : heads * dv].reshape(nb, L, heads, dv).permute(2, 0, 1, 3)
    o = (a @ v).permute(1, 2, 0, 3).reshape(nb, L, heads * dv)
    return o * gate[..., : heads * dv] if gate is not None else o


@pytest.mark.parametrize("L", [1, 37, 64, 130, 611])
@pytest.mark.parametrize("kind", ["self", "self_mma", "nonlin384", "nonlin144", "nonlin384_mma", "nonlin144_mma",
                                  "nonlin384_redo", "nonlin144_redo"])
def test_attention(L, kind):
    H, nb = 4, 3
    # _mma: the mma.sync flash kernel; _redo: the tcgen05 NonlinAttention kernel with its exact-maxima repeat forced
    _lib().zv_debug_set_attn_impl(1 if kind.endswith("_mma") else 2 if kind.endswith("_redo") else 0)
    kind = kind.replace("_mma", "").replace("_redo", "")
    g = torch.Generator(device="cuda").manual_seed(L)
    rn = lambda *s: torch.randn(*s, device="cuda", generator=g)
    qkp = (rn(nb, L, 288) * 0.7).bfloat16().float()
    E = rn(H, 2 * L - 1, 4).contiguous()
    lens = torch.tensor([L, max(1, L // 2), max(1, L - 3)], device="cuda")
    mask = (torch.arange(L, device="cuda")[None] >= lens[:, None]).to(torch.uint8).contiguous()
    heads, dv = (4, 12) if kind == "self" else (1, int(kind[6:]))
    ldv = heads * dv
    V = rn(nb, L, ldv).bfloat16().float()
    gate = None if kind == "self" else rn(nb, L, ldv).bfloat16().float()
    out = torch.zeros(nb, L, ldv, device="cuda")
    rc = _lib().zv_debug_attention(nb, L, H, heads, dv, _p(qkp), 288, _p(E), _p(mask), _p(V), ldv, _p(gate), _p(out),
                                   None)
    _lib().zv_debug_set_attn_impl(0)
    assert rc == 0
    want = ref_attention(qkp, E, mask, V, gate, H, heads, dv)
    r = rel(out, want)
    assert r < 8e-3, f"rel-L2 {r:.3e} max-abs {float((out - want).abs().max()):.3e}"


@pytest.mark.parametrize("L", [130, 611])
def test_nonlin_attention_overflow_repeat(L):
    """Scores of late keys sit ~130 nats above those of key tile 0 and of the early diagonal tiles: the optimistic
    value sweep of the tcgen05 NonlinAttention kernel must notice and repeat with exact row maxima."""
    H, nb, dv = 4, 2, 384
    g = torch.Generator(device="cuda").manual_seed(L)
    rn = lambda *s: torch.randn(*s, device="cuda", generator=g)
    qkp = (rn(nb, L, 288) * 0.05).bfloat16().float()
    qkp[..., :32] += 2.0                                   # q of head 0
    qkp[:, L // 2:, H * 32: H * 32 + 32] += 2.0            # k of head 0, second half of the keys: s ~ 128
    qkp = qkp.bfloat16().float()
    E = rn(H, 2 * L - 1, 4).contiguous()
    mask = torch.zeros(nb, L, dtype=torch.uint8, device="cuda")
    V = rn(nb, L, dv).bfloat16().float()
    gate = rn(nb, L, dv).bfloat16().float()
    out = torch.zeros(nb, L, dv, device="cuda")
    rc = _lib().zv_debug_attention(nb, L, H, 1, dv, _p(qkp), 288, _p(E), _p(mask), _p(V), dv, _p(gate), _p(out), None)
    assert rc == 0
    want = ref_attention(qkp, E, mask, V, gate, H, 1, dv)
    assert torch.isfinite(out).all()
    r = rel(out, want)
    assert r < 8e-3, f"rel-L2 {r:.3e}"


@pytest.mark.parametrize("L", [130, 611, 1219])
def test_self_attention_reference_repeat(L):
    """SelfAttention kernel: interior key tiles are exponentiated against the running reference maximum without
    waiting for their own maximum.  Scores of the late keys sit ~130 nats above everything before them, far over the
    2^64 bound: the warp must notice, repeat that tile with the exact online-softmax step and carry on."""
    H, nb, dv = 4, 2, 12
    g = torch.Generator(device="cuda").manual_seed(L)
    rn = lambda *s: torch.randn(*s, device="cuda", generator=g)
    qkp = (rn(nb, L, 288) * 0.05).bfloat16().float()
    qkp[..., : H * 32] += 2.0                              # q of every head
    qkp[:, L // 2:, H * 32: 2 * H * 32] += 2.0             # k of every head, second half of the keys: s ~ 128
    qkp = qkp.bfloat16().float()
    E = rn(H, 2 * L - 1, 4).contiguous()
    mask = torch.zeros(nb, L, dtype=torch.uint8, device="cuda")
    mask[1, L - 5:] = 1
    V = rn(nb, L, H * dv).bfloat16().float()
    out = torch.zeros(nb, L, H * dv, device="cuda")
    rc = _lib().zv_debug_attention(nb, L, H, H, dv, _p(qkp), 288, _p(E), _p(mask), _p(V), H * dv, None, _p(out), None)
    assert rc == 0
    want = ref_attention(qkp, E, mask, V, None, H, H, dv)
    assert torch.isfinite(out).all()
    r = rel(out, want)
    assert r < 8e-3, f"rel-L2 {r:.3e}"


@pytest.mark.parametrize("impl", [1, 2, 3], ids=["simt", "tcgen05", "tcgen05_pair"])
def test_gemm_fp16_output_is_precise_at_1e3_and_saturates(impl):
    """The attention q | k | p projection stores IEEE fp16, un-scaled.  At |q| ~ 1e3 the stored value
    keeps fp16's 11-bit mantissa; beyond the fp16 range it SATURATES at +-65504 (cvt.rn.satfinite) instead of turning
    into +-inf, which would make a whole softmax row NaN."""
    g = torch.Generator(device="cuda").manual_seed(5)
    M, N, K = 300, 272, 512   # the q | k | p projection's shape
    A = torch.randn(M, K, device="cuda", generator=g)
    W = torch.randn(N, K, device="cuda", generator=g) * 0.05
    bias = torch.tensor([0.0, 1e3, -1e3, 1e5, -1e5, 6.6e4, 3e38], device="cuda")[torch.arange(N, device="cuda") % 7].contiguous()
    out, _ = run_gemm(7, impl, A, W, bias)          # mode 7: EPI_BIAS with the fp16 store
    ref, _ = ref_gemm(EPI_BIAS, A, W, bias)
    assert bool(torch.isfinite(out).all()), "fp16 overflow to inf"
    want = ref.clamp(-65504.0, 65504.0)
    err = (out - want).abs()
    assert bool((err <= want.abs() * 2.0 ** -11 + 1e-3).all()), float((err / (want.abs() + 1.0)).max())
    big = (torch.arange(N, device="cuda") % 7 >= 3)
    assert bool((out[:, big].abs() == 65504.0).all())
    k3 = (torch.arange(N, device="cuda") % 7 == 1)
    assert float((out[:, k3] - ref[:, k3]).abs().max()) <= 0.5 + 1e-3   # spacing of fp16 near 1e3 is 0.5


@pytest.mark.parametrize("ksplit", [2, 3, 8])
@pytest.mark.parametrize("L", [130, 611, 1219])
def test_self_attention_key_split(L, ksplit, monkeypatch):
    """The SelfAttention kernel's key split (small grids: single utterances, the 60 s dialog): ksplit CTAs per query
    tile write partial softmax states that a merge kernel combines.  Forced here (the launcher would pick it from the
    grid size); must equal the unsplit kernel to rounding and the fp32 reference to the attention tolerance."""
    H, nb, dv = 4, 2, 12
    g = torch.Generator(device="cuda").manual_seed(L + ksplit)
    rn = lambda *s: torch.randn(*s, device="cuda", generator=g)
    qkp = (rn(nb, L, 288) * 0.7).bfloat16().float()
    qkp[:, L // 2:, H * 32: 2 * H * 32] += 1.0      # later keys score higher: the splits end with different maxima
    qkp = qkp.bfloat16().float()
    E = rn(H, 2 * L - 1, 4).contiguous()
    mask = torch.zeros(nb, L, dtype=torch.uint8, device="cuda")
    mask[1, L - 7:] = 1
    V = rn(nb, L, H * dv).bfloat16().float()
    outs = {}
    for k in (1, ksplit):
        out = torch.zeros(nb, L, H * dv, device="cuda")
        rc = _lib().zv_debug_attention_ksplit(k)
        assert rc == 0
        rc = _lib().zv_debug_attention(nb, L, H, H, dv, _p(qkp), 288, _p(E), _p(mask), _p(V), H * dv, None, _p(out), None)
        assert rc == 0
        outs[k] = out
    _lib().zv_debug_attention_ksplit(0)
    want = ref_attention(qkp, E, mask, V, None, H, H, dv)
    assert rel(outs[ksplit], want) < 8e-3
    d = float((outs[ksplit] - outs[1]).abs().max())
    assert d < 2e-2 and rel(outs[ksplit], outs[1]) < 4e-3, (d, rel(outs[ksplit], outs[1]))
    nt = (L + 63) // 64
    per = (nt + ksplit - 1) // ksplit
    if per >= 2 and (ksplit - 1) * per < nt:   # a legal split (no empty key range) was really taken: not bit-identical
        assert not torch.equal(outs[ksplit], outs[1])

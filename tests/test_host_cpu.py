"""CPU-only checks of the host side: the C-ABI library loads and exports every symbol declared in
include/zipvoice_b200.h, pure-host entry points agree with torch, and the product fails loudly
without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from tests.conftest import ROOT


def test_library_exports_every_declared_symbol():
    from zipvoice_b200 import _lib
    lib = _lib.load()
    hdr = open(os.path.join(ROOT, "include", "zipvoice_b200.h")).read()
    declared = set(re.findall(r"\b(zv_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/zipvoice_b200.h but not exported"
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    assert not any(n.startswith("zv_debug") for n in declared), "test hooks belong in csrc/zv_debug.h"
    dbg = open(os.path.join(ROOT, "zipvoice_b200", "csrc", "zv_debug.h")).read()
    private = set(re.findall(r"\b(zv_[a-z0-9_]+)\s*\(", dbg))
    assert private == set(_lib.DEBUG_EXPORTS), private ^ set(_lib.DEBUG_EXPORTS)
    for name in sorted(private):
        assert hasattr(lib, name), f"{name} declared in csrc/zv_debug.h but not exported"


def test_time_steps_match_torch():
    from oracle import zipvoice_oracle as O
    from zipvoice_b200.solver import get_time_steps
    for n, s, a, b in [(16, 0.5, 0.0, 1.0), (4, 0.7, 0.0, 1.0), (7, 1.0, 0.2, 0.9), (1, 0.5, 0.0, 1.0),
                       (8, 0.3, 0.0, 1.0), (33, 0.9, 0.1, 0.7)]:
        assert torch.equal(get_time_steps(a, b, n, s), O.get_time_steps(a, b, n, s)), (n, s, a, b)


def test_config_struct_matches_header():
    from zipvoice_b200 import _lib, synth
    from zipvoice_b200.engine import make_config
    c = make_config(synth.BASE_MODEL_CFG, synth.VOCAB_SIZE, synth.PAD_ID, distill=True)
    assert list(c.fm_downsampling_factor)[:5] == [1, 2, 4, 2, 1] and c.fm_num_stacks == 5
    assert c.use_guidance_scale_embed == 1 and c.vocab_size == 364
    hdr = open(os.path.join(ROOT, "include", "zipvoice_b200.h")).read()
    body = hdr[hdr.index("typedef struct zv_config {"):hdr.index("} zv_config;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = re.findall(r"\b([a-z_0-9]+)(?:\[[A-Z_]+\])?\s*[,;]", body)
    assert names == [f[0] for f in _lib.ZvConfig._fields_], names


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    from zipvoice_b200 import synth
    from zipvoice_b200.engine import Engine, ZvError
    with pytest.raises(ZvError):
        Engine(synth.small_model_cfg(), {}, synth.VOCAB_SIZE)
    from zipvoice_b200 import _lib
    h = C.c_void_p()
    cfg = _lib.ZvConfig()
    assert _lib.load().zv_create(C.byref(cfg), C.byref(h)) != 0


def test_synth_state_dict_is_deterministic():
    from zipvoice_b200 import synth
    a = synth.make_state_dict(synth.small_model_cfg(), seed=3, stress=True)
    b = synth.make_state_dict(synth.small_model_cfg(), seed=3, stress=True)
    assert a.keys() == b.keys() and all(torch.equal(a[k], b[k]) for k in a)
    full = synth.make_state_dict(distill=True)
    assert len(full) == 890 and abs(sum(v.numel() for v in full.values()) / 1e6 - 122.7) < 0.1


def test_s24_stream_format_restated_in_numpy():
    """The residual stream's storage format (csrc/zv_common.cuh: bf16 `hi` plane + byte `lo` plane): restated here in
    numpy integer arithmetic and checked for what the kernels rely on -- decode(encode(v)) is v rounded to 24 bits
    (relative error <= 2^-16), and `hi` alone is a correctly rounded bf16 of v (half up instead of half to even: it
    differs from torch's bf16 by at most one bf16 ulp and only where the discarded bits sit within 2^-8 of a tie).
    The CUDA side is checked against the same fp32 references through the GEMM tests (`EPI_RES24` cases)."""
    rng = np.random.default_rng(24)
    v = np.concatenate([rng.standard_normal(200000).astype(np.float32) * 3.0,
                        (rng.standard_normal(50000) * 1e-6).astype(np.float32),
                        (rng.standard_normal(50000) * 1e6).astype(np.float32),
                        np.array([0.0, -0.0, 1.0, -1.0, 65504.0, 2.0 ** -126, np.inf, -np.inf], np.float32)])
    bits = v.view(np.uint32)
    w = bits + np.uint32(0x8080)
    hi = (w >> np.uint32(16)).astype(np.uint16)
    lo = ((w >> np.uint32(8)) & np.uint32(0xff)).astype(np.uint8)
    dec = (((hi.astype(np.uint32) << np.uint32(16)) | (lo.astype(np.uint32) << np.uint32(8))) - np.uint32(0x8000)).view(np.float32)
    fin = np.isfinite(v)
    assert np.array_equal(dec[~fin], v[~fin])
    err = np.abs(dec[fin].astype(np.float64) - v[fin].astype(np.float64))
    assert np.all(err <= np.abs(v[fin]).astype(np.float64) * 2.0 ** -16)
    assert np.array_equal(dec.view(np.uint32) & np.uint32(0xff), np.zeros_like(bits))  # 24 bits: the low byte is clear
    # re-encoding a decoded value is the identity (in-place updates of unchanged elements do not drift)
    w2 = dec.view(np.uint32) + np.uint32(0x8080)
    assert np.array_equal((w2 >> np.uint32(16)).astype(np.uint16), hi) and \
        np.array_equal(((w2 >> np.uint32(8)) & np.uint32(0xff)).astype(np.uint8), lo)
    # hi against torch's round-to-nearest-even bf16
    tb = torch.from_numpy(v[fin]).bfloat16().view(torch.int16).numpy().view(np.uint16)
    d = hi[fin].astype(np.int64) - tb.astype(np.int64)
    assert np.all(np.abs(d) <= 1)
    low16 = bits[fin] & np.uint32(0xffff)
    assert np.all((d == 0) | ((low16 >= 0x7f80) & (low16 <= 0x8000))), "hi differs from bf16 RNE away from a tie"

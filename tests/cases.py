"""Seeded solver-level inputs shared by oracle/make_golden.py and the tests."""
import torch


def case_inputs(B, T, prompt, seed, ragged=True):
    g = torch.Generator().manual_seed(seed)
    x0 = torch.randn(B, T, 100, generator=g)
    text = torch.randn(B, T, 100, generator=g) * 0.5
    speech = torch.randn(B, T, 100, generator=g) * 0.5 - 0.5
    speech[:, prompt:] = 0
    lens = torch.full((B,), T, dtype=torch.int64)
    if ragged and B > 1:
        lens[1:] = torch.randint(T // 2, T, (B - 1,), generator=g)
    mask = torch.arange(T)[None] >= lens[:, None]
    return x0, text, speech, mask


def case_inputs_stereo(B, T, prompt, seed):
    """Two-channel dialog model (ZipVoiceDialogStereo): x / speech are 200 wide, the text condition 100."""
    g = torch.Generator().manual_seed(seed)
    x0 = torch.randn(B, T, 200, generator=g)
    text = torch.randn(B, T, 100, generator=g) * 0.5
    speech = torch.randn(B, T, 200, generator=g) * 0.5 - 0.5
    speech[:, prompt:] = 0
    lens = torch.full((B,), T, dtype=torch.int64)
    if B > 1:
        lens[1:] = torch.randint(T // 2, T, (B - 1,), generator=g)
    mask = torch.arange(T)[None] >= lens[:, None]
    return x0, text, speech, mask


def rel_l2(a, b):
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def valid_rel_l2(a, b, mask):
    keep = (~mask).unsqueeze(-1).to(a.dtype)
    return rel_l2(a * keep, b * keep)


def headline_case(name):
    """Full-size solver-level inputs (BASELINE.json configs 2 / 3 / 4 at B = 2 / 2 / 1) shared by
    oracle/make_golden.py headline and tests/test_gpu_headline.py."""
    if name == "zipvoice":       # config 2 shape: T = 1219 = 281-frame prompt + 938 generated frames
        return case_inputs(2, 1219, 281, seed=41)
    if name == "distill":        # config 3 shape
        return case_inputs(2, 1219, 281, seed=42)
    if name == "dialog":         # config 4 shape: 10 s prompt (938 frames) + 60 s (5625 frames)
        return case_inputs(1, 6563, 938, seed=43)
    if name == "stress":
        return case_inputs(1, 1219, 281, seed=44)
    raise KeyError(name)


def guided_tol(g_eff, tol_raw):
    """Per-NFE tolerance of a classifier-free-guided velocity v = (1 + g) c - g u (solver.py:100-110): the two network
    passes carry independent rounding errors of relative size tol_raw and |v| ~ |c| ~ |u|, so the guided bound is
    tol_raw * sqrt((1 + g)^2 + g^2); g_eff is the guidance actually applied (doubled for t <= 0.5, solver.py:92-97)."""
    return tol_raw * ((1.0 + g_eff) ** 2 + g_eff ** 2) ** 0.5

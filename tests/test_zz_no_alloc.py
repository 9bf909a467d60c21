"""zv_resample_static: the allocation-free variant of zv_resample (include/zipvoice_b200.h; no counterpart in the
reference) against the existing entry point, through the C ABI.

The file is named so that pytest collects it LAST: a failure here hides nothing.
It builds its own handle and calls the static entry point BEFORE zv_resample has seen the rate pair, so it passes
alone in a fresh process as well as after the other tests.  Tolerance: the 2e-5 absolute of the resampler tests
(tests/test_gpu_audio.py) on |x| <= 1 signals; the two entry points share the resampling kernel and differ only in
where the fp32 filter bank was computed (device double vs host double arithmetic)."""
import os

import numpy as np
import pytest
import torch

from tests.conftest import GOLDEN

pytestmark = pytest.mark.gpu

TOL = 2e-5


@pytest.fixture(scope="module")
def eng():
    from zipvoice_b200.audio import audio_engine
    e = audio_engine(device="cuda:0")
    yield e
    e.close()


def _n_out(n, orig, new):
    return (n * new + orig - 1) // orig


def test_resample_static_two_calls_in_a_row_match_resample(eng):
    dev = eng.device
    g = torch.Generator().manual_seed(17)
    x = (torch.randn(3, 7000, generator=g) * 0.3).to(dev)
    lens = torch.tensor([7000, 4411, 1], dtype=torch.int64, device=dev)
    # every buffer the static entry point touches is allocated here, by the caller, up front
    calls = [(44100, 24000, lens), (16000, 24000, None), (44100, 24000, lens), (22050, 24000, None)]
    outs = [torch.full((3, _n_out(7000, o, n)), 7.0, device=dev) for o, n, _ in calls]
    torch.cuda.synchronize()
    got = []
    for i, ((o, n, ln), out) in enumerate(zip(calls, outs)):
        if i >= 1:  # call 0 may load the kernels' module; from call 1 on the device's free memory must not move
            free0, alloc0 = torch.cuda.mem_get_info(dev)[0], torch.cuda.memory_allocated(dev)
        r = eng.resample_static(x, o, n, out, ln)
        assert r is out
        if i >= 1:
            torch.cuda.synchronize()
            assert torch.cuda.mem_get_info(dev)[0] == free0, "zv_resample_static changed the device's free memory"
            assert torch.cuda.memory_allocated(dev) == alloc0
        got.append(out.cpu())
    # the existing (allocating) entry point, called only now: the static path cannot have leaned on its filter banks
    for i, ((o, n, ln), mine) in enumerate(zip(calls, got)):
        want = eng.resample(x, o, n, ln).cpu()
        assert mine.shape == want.shape
        d = float((mine - want).abs().max())
        assert d < TOL, f"call {i} ({o} -> {n} Hz): max-abs {d:.3e} vs zv_resample"
        assert float(want.abs().max()) > 1e-2   # not comparing zeros with zeros
    # ragged rows end where zv_resample ends them (rows past ceil(new * len / orig) are exactly 0)
    assert float(got[0][2, _n_out(1, 44100, 24000):].abs().max()) == 0.0
    # calls 0 and 2 bracket a call with another rate pair: the shared scratch was rebuilt, results are bit-identical
    assert torch.equal(got[0], got[2])


def test_resample_static_matches_torchaudio_golden(eng):
    """Two calls in a row on the golden inputs of the existing resampler test (tests/golden/resample.npz)."""
    g = np.load(os.path.join(GOLDEN, "resample.npz"))
    for sr in (48000, 44100):
        x = torch.from_numpy(g[f"x_{sr}"]).to(eng.device).contiguous()
        want = torch.from_numpy(g[f"y_{sr}"])
        out = torch.empty(x.shape[0], _n_out(x.shape[1], sr, 24000), device=eng.device)
        eng.resample_static(x, sr, 24000, out)
        assert out.shape == want.shape
        assert float((out.cpu() - want).abs().max()) < TOL
        assert float((out.cpu() - eng.resample(x, sr, 24000).cpu()).abs().max()) < TOL


def test_resample_static_rejects_banks_beyond_the_static_array(eng):
    from zipvoice_b200.engine import ZvError
    x = torch.zeros(1, 5000, device=eng.device)
    out = torch.full((1, _n_out(5000, 44101, 24000)), 3.0, device=eng.device)
    with pytest.raises(ZvError) as ei:   # gcd(44101, 24000) = 1: a bank of 24000 phases x 44125 taps
        eng.resample_static(x, 44101, 24000, out)
    assert "ZV_RESAMPLE_STATIC_MAX_FLOATS" in str(ei.value) and "967" in str(ei.value)
    torch.cuda.synchronize()
    assert bool((out == 3.0).all()), "a rejected call must not launch anything"
    assert eng.resample(x, 44100, 24000).shape == (1, _n_out(5000, 44100, 24000))  # the handle is still usable

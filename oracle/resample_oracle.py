"""TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline): CPU restatement of the
prompt resampling step of the reference.

The reference calls ``torchaudio.transforms.Resample(orig_freq, new_freq)`` on the prompt recording
(zipvoice/bin/infer_zipvoice.py:334-339, bin/infer_zipvoice_dialog.py prompt loading).  torchaudio is a third-party
dependency that is not vendored under /root/reference (requirements.txt: ``torchaudio``, unpinned; 2.11.0 is the
version in this image); its published algorithm (torchaudio/functional/functional.py,
``_get_sinc_resample_kernel`` / ``_apply_sinc_resample_kernel`` with the defaults sinc_interp_hann,
lowpass_filter_width = 6, rolloff = 0.99) is restated here with numpy.  Pinned by tests/golden/resample.npz, which
``python -m oracle.make_golden resample`` generates with torchaudio itself.
"""
import math

import numpy as np


def sinc_kernel(orig_freq: int, new_freq: int, lowpass_filter_width: int = 6, rolloff: float = 0.99):
    """-> (kernel (new, K) float32, width, orig, new) with the frequencies reduced by their gcd."""
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    base_freq = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / base_freq)
    idx = np.arange(-width, width + orig, dtype=np.float64)[None, :] / orig
    t = np.arange(0, -new, -1, dtype=np.float64)[:, None] / new + idx
    t = t * base_freq
    t = np.clip(t, -lowpass_filter_width, lowpass_filter_width)
    window = np.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t = t * math.pi
    scale = base_freq / orig
    with np.errstate(invalid="ignore", divide="ignore"):
        kern = np.where(t == 0, 1.0, np.sin(t) / t)
    kern = kern * (window * scale)
    return kern.astype(np.float32), width, orig, new


def resample(wav: np.ndarray, orig_freq: int, new_freq: int) -> np.ndarray:
    """wav (..., N) float32 -> (..., ceil(new * N / orig)) float32."""
    wav = np.asarray(wav, dtype=np.float32)
    if orig_freq == new_freq:
        return wav
    kern, width, orig, new = sinc_kernel(orig_freq, new_freq)
    lead = wav.shape[:-1]
    x = wav.reshape(-1, wav.shape[-1])
    n = x.shape[-1]
    xp = np.pad(x, ((0, 0), (width, width + orig)))
    K = kern.shape[1]
    n_frames = (xp.shape[-1] - K) // orig + 1
    # frames[b, k, j] = xp[b, k * orig + j]
    sb, sn = xp.strides
    frames = np.lib.stride_tricks.as_strided(xp, (x.shape[0], n_frames, K), (sb, sn * orig, sn), writeable=False)
    out = np.einsum("bkj,pj->bkp", frames, kern, optimize=True).astype(np.float32).reshape(x.shape[0], -1)
    target = -(-new * n // orig)
    return out[:, :target].reshape(*lead, target)

"""TEST INFRASTRUCTURE ONLY -- numpy/torch CPU restatement of VocosFbank
(zipvoice/utils/feature.py:36-120): torchaudio MelSpectrogram(sample_rate 24000, n_fft 1024,
win 1024 periodic Hann, hop 256, center=True reflect padding, power 1, 100 HTK mel bins,
f 0..12000, norm None) -> clamp(1e-7) -> log -> (T, 100), T forced to lhotse's
compute_num_frames.

Parity PINNED for everything up to the frame count: tests/golden/fbank_*.npz holds the output of
the unmodified reference class (imported with a minimal ``lhotse`` shim, oracle/make_golden.py).
The frame count itself follows lhotse.utils.compute_num_frames (third-party, unpinned
requirements.txt:6, source absent): round-half-up of duration / frame_shift = (N + 128) // 256;
UNPINNED at that one boundary (differs from torchaudio's 1 + N // 256 only when N % 256 < 128).
"""
import math

import numpy as np
import torch

SR, N_FFT, HOP, N_MELS = 24000, 1024, 256, 100


def num_frames(n_samples: int) -> int:
    """lhotse compute_num_frames(duration, frame_shift, sr) with ROUND_HALF_UP (feature.py:100-102)."""
    return (n_samples + HOP // 2) // HOP


def mel_filterbank() -> np.ndarray:
    """torchaudio.functional.melscale_fbanks(513, 0, 12000, 100, 24000, norm=None, mel_scale='htk')
    -> (513, 100), evaluated in float64 and rounded to fp32."""
    n_freqs = N_FFT // 2 + 1
    all_freqs = np.linspace(0.0, SR // 2, n_freqs)
    hz2mel = lambda f: 2595.0 * math.log10(1.0 + f / 700.0)
    m_pts = np.linspace(hz2mel(0.0), hz2mel(SR / 2.0), N_MELS + 2)
    f_pts = 700.0 * (10.0 ** (m_pts / 2595.0) - 1.0)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts[None, :] - all_freqs[:, None]
    down = -slopes[:, :-2] / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    return np.maximum(0.0, np.minimum(down, up)).astype(np.float32)


def vocos_fbank(samples: torch.Tensor) -> torch.Tensor:
    """samples (N,) or (1, N) fp32 mono 24 kHz -> log-mel (T, 100) (feature.py:56-116, num_channels=1)."""
    x = samples.reshape(-1).to(torch.float32)
    n = x.numel()
    pad = N_FFT // 2
    xp = torch.nn.functional.pad(x[None, None], (pad, pad), mode="reflect")[0, 0]
    frames = xp.unfold(0, N_FFT, HOP)                      # (1 + N // 256, 1024)
    win = torch.hann_window(N_FFT, periodic=True)
    spec = torch.fft.rfft(frames * win, dim=-1).abs()      # power = 1
    mel = spec @ torch.from_numpy(mel_filterbank())
    logmel = mel.clamp(min=1e-7).log()
    t = num_frames(n)
    if logmel.shape[0] > t:
        logmel = logmel[:t]
    elif logmel.shape[0] < t:
        logmel = torch.cat([logmel, logmel[-1:].expand(t - logmel.shape[0], -1)], 0)
    return logmel


# ---------------------------------------------------------------- BigVGAN features (alternative vocoder config)
def bigvgan_mel_basis() -> np.ndarray:
    """librosa.filters.mel(sr=24000, n_fft=1024, n_mels=100, fmin=0, fmax=None) -> (100, 513), Slaney scale and
    area normalisation (zipvoice/utils/_bigvgan_mel_feature.py:76-79; restated in oracle/shims/librosa/filters.py)."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "shims"))
    try:
        from librosa.filters import mel
    finally:
        sys.path.pop(0)
    return mel(sr=SR, n_fft=N_FFT, n_mels=N_MELS, fmin=0, fmax=None)


def bigvgan_fbank(samples: torch.Tensor) -> torch.Tensor:
    """BigVGANFbank.extract, num_channels=1 (zipvoice/utils/feature.py:133-204 on
    _bigvgan_mel_feature.py:42-111): reflect padding (1024 - 256) / 2 = 384 per side, frames WITHOUT centering,
    periodic Hann, sqrt(re^2 + im^2 + 1e-9), Slaney mel basis, log(clamp(., 1e-5)) -> (T, 100) with T forced to
    lhotse's compute_num_frames (replicating the last frame: the un-centred STFT yields N // 256 frames)."""
    x = samples.reshape(-1).to(torch.float32)
    n = x.numel()
    pad = (N_FFT - HOP) // 2
    xp = torch.nn.functional.pad(x[None, None], (pad, pad), mode="reflect")[0, 0]
    frames = xp.unfold(0, N_FFT, HOP)
    win = torch.hann_window(N_FFT, periodic=True)
    spec = torch.fft.rfft(frames * win, dim=-1)
    mag = torch.sqrt(spec.real ** 2 + spec.imag ** 2 + 1e-9)
    mel = mag @ torch.from_numpy(bigvgan_mel_basis()).T
    logmel = torch.log(torch.clamp(mel, min=1e-5))
    t = num_frames(n)
    if logmel.shape[0] > t:
        logmel = logmel[:t]
    elif logmel.shape[0] < t:
        logmel = torch.cat([logmel, logmel[-1:].expand(t - logmel.shape[0], -1)], 0)
    return logmel

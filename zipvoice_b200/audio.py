"""Drop-in objects for the two audio seams of ``generate_sentence(model, vocoder, tokenizer,
feature_extractor, ...)`` (zipvoice/bin/infer_zipvoice.py:276-294):

* ``VocosFbankB200.extract(samples, sampling_rate)``  mirrors ``VocosFbank.extract``
  (zipvoice/utils/feature.py:63-116);
* ``VocosB200.decode(mel)``  mirrors ``vocos.Vocos.decode`` (call site infer_zipvoice.py:378).

Both run entirely in the CUDA library; there is no CPU path.
"""
from typing import Dict, Optional, Union

import numpy as np
import torch

from . import synth
from .engine import Engine


def audio_engine(vocoder_state_dict: Optional[Dict[str, torch.Tensor]] = None, device="cuda:0") -> Engine:
    """A handle that only carries the mel front-end and (optionally) the vocoder weights."""
    sd = {}
    if vocoder_state_dict is not None:
        sd = {("vocoder." + k): v for k, v in vocoder_state_dict.items()
              if k.startswith(("backbone.", "head.out."))}
    return Engine(synth.BASE_MODEL_CFG, sd, synth.VOCAB_SIZE, synth.PAD_ID, device=device)


class VocosFbankB200:
    """num_channels 1 (mono; stereo input is averaged) or 2 (per-channel mels concatenated)."""

    sampling_rate, n_mels, n_fft, hop_length = 24000, 100, 1024, 256
    kind = "vocos"  # which feature kernel of the engine (BigVGANFbankB200 overrides)

    def __init__(self, engine: Optional[Engine] = None, num_channels: int = 1, device="cuda:0"):
        assert num_channels in (1, 2)
        self.num_channels = num_channels
        self.engine = engine if engine is not None else audio_engine(device=device)

    @property
    def frame_shift(self) -> float:
        return self.hop_length / self.sampling_rate

    def feature_dim(self, sampling_rate: int) -> int:
        return self.n_mels

    def extract(self, samples: Union[np.ndarray, torch.Tensor], sampling_rate: int):
        assert sampling_rate == self.sampling_rate, (
            f"Mismatched sampling rate: extractor expects {self.sampling_rate}, got {sampling_rate}")
        is_numpy = not isinstance(samples, torch.Tensor)
        if is_numpy:
            samples = torch.from_numpy(samples)
        if samples.dim() == 1:
            samples = samples.unsqueeze(0)
        else:
            assert samples.ndim == 2, samples.shape
        if self.num_channels == 1:
            if samples.shape[0] == 2:
                samples = samples.mean(dim=0, keepdims=True)
        else:
            assert samples.shape[0] == 2, samples.shape
        mel = self.engine.fbank(samples.float(), kind=self.kind)  # (C, T, 100)
        mel = mel.permute(1, 0, 2).reshape(mel.shape[1], -1)     # (T, 100 * C)
        return mel.cpu().numpy() if is_numpy else mel

    def extract_batch(self, samples: torch.Tensor, lens: Optional[torch.Tensor] = None) -> torch.Tensor:
        """(B, N) padded mono batch with valid lengths -> (B, T, 100); frames past an utterance are 0."""
        return self.engine.fbank(samples, lens, kind=self.kind)


class BigVGANFbankB200(VocosFbankB200):
    """BigVGANFbank (zipvoice/utils/feature.py:133-204; feature type "bigvgan_v2", bin/infer_zipvoice.py:585-586): same
    interface and frame count as VocosFbank, features = log(clamp(slaney_mel(sqrt(|STFT|^2 + 1e-9)), 1e-5)) of frames
    that are NOT centred (reflect padding (n_fft - hop) / 2).  The bigvgan vocoder itself is a third-party network
    (``bigvgan`` package, absent here) and is not part of this path."""

    kind = "bigvgan"


class VocosB200:
    def __init__(self, state_dict: Dict[str, torch.Tensor], engine: Optional[Engine] = None, device="cuda:0"):
        self.engine = engine if engine is not None else audio_engine(state_dict, device=device)

    def eval(self):
        return self

    def to(self, device):
        return self

    @torch.inference_mode()
    def decode(self, features_input: torch.Tensor, lens: Optional[torch.Tensor] = None) -> torch.Tensor:
        """mel (B, 100, T) -> (B, 256 (T - 1))"""
        return self.engine.vocos_decode(features_input, lens)


class ResampleB200:
    """torchaudio.transforms.Resample(orig_freq, new_freq) with torchaudio's defaults, on the GPU
    (call site: bin/infer_zipvoice.py:334-339)."""

    def __init__(self, orig_freq: int, new_freq: int, engine: Optional[Engine] = None, device="cuda:0"):
        self.orig_freq, self.new_freq = int(orig_freq), int(new_freq)
        self.engine = engine if engine is not None else audio_engine(device=device)

    def __call__(self, waveform: torch.Tensor, lens: Optional[torch.Tensor] = None) -> torch.Tensor:
        if self.orig_freq == self.new_freq:
            return waveform
        shape = waveform.shape
        out = self.engine.resample(waveform.reshape(-1, shape[-1]).float(), self.orig_freq, self.new_freq, lens)
        return out.reshape(*shape[:-1], out.shape[-1])


def compute_fbank(waveforms, extractor: Optional[VocosFbankB200] = None, max_batch_samples: int = 64 * 30 * 24000,
                  out_path: Optional[str] = None):
    """Bulk offline feature extraction on the mel kernel (bin/compute_fbank.py:136-260 without lhotse: the caller
    hands over 24 kHz mono recordings, the result is a headless buffer instead of the Lilcom writer).

    waveforms: sequence of 1-D float arrays / tensors.  They are sorted by length, packed into padded batches of
    at most ``max_batch_samples`` samples, extracted with per-row lengths, and returned in the input order as a
    list of (T_i, 100) float32 numpy arrays (T_i = (n_i + 128) // 256).  With ``out_path`` the features are also
    written as one ``.npz`` (arrays ``feat_<index>``)."""
    ext = extractor if extractor is not None else VocosFbankB200()
    dev = ext.engine.device
    lens = [int(len(w)) for w in waveforms]
    order = sorted(range(len(lens)), key=lambda i: lens[i])
    feats = [None] * len(lens)
    i = 0
    while i < len(order):
        j, longest = i, 0
        while j < len(order) and max(longest, lens[order[j]]) * (j - i + 1) <= max(max_batch_samples, lens[order[i]]):
            longest = max(longest, lens[order[j]])
            j += 1
        idx = order[i:j]
        batch = torch.zeros(len(idx), longest, dtype=torch.float32)
        for r, k in enumerate(idx):
            batch[r, : lens[k]] = torch.as_tensor(waveforms[k], dtype=torch.float32).reshape(-1)
        ln = torch.tensor([lens[k] for k in idx], dtype=torch.int64)
        mel = ext.extract_batch(batch.to(dev, non_blocking=True), ln.to(dev)).cpu().numpy()
        for r, k in enumerate(idx):
            feats[k] = mel[r, : (lens[k] + 128) // 256].copy()
        i = j
    if out_path is not None:
        np.savez(out_path, **{f"feat_{k}": f for k, f in enumerate(feats)})
    return feats

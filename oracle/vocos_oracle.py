"""TEST INFRASTRUCTURE ONLY -- torch-fp32 CPU restatement of ``Vocos.decode`` for the
``charactr/vocos-mel-24khz`` configuration (call sites: zipvoice/bin/infer_zipvoice.py:249-260,378;
zipvoice/bin/infer_zipvoice_dialog.py:192-203,322).

PARITY UNPINNED: ``vocos`` is a third-party dependency of the reference (requirements.txt:10,
unpinned), its source and weights are not under /root/reference and the reference has no test or
fixture for it.  This file restates the published architecture (VocosBackbone + ISTFTHead):
  backbone.embed  Conv1d(100, 512, k=7, pad=3) -> LayerNorm(512, eps 1e-6)
  8 x ConvNeXtBlock: dwconv Conv1d(512, 512, k=7, pad=3, groups=512) -> LayerNorm(eps 1e-6)
      -> Linear(512, 1536) -> GELU (erf) -> Linear(1536, 512) -> gamma * -> + residual
  final LayerNorm(eps 1e-6) -> head.out Linear(512, 1026) -> mag = exp(.)clip(max 1e2), phase
  -> S = mag (cos p + i sin p) -> torch.istft(n_fft 1024, hop 256, win 1024 Hann, center=True).
State-dict names follow the vocos checkpoint (``backbone.*``, ``head.out.*``).
"""
from typing import Dict

import torch
import torch.nn.functional as F

N_FFT, HOP, DIM, INTER, LAYERS, N_MELS = 1024, 256, 512, 1536, 8, 100


def make_vocos_state_dict(seed: int = 0) -> Dict[str, torch.Tensor]:
    """Seeded random vocoder weights (shared with the product-side synthetic generator)."""
    from zipvoice_b200.synth import make_vocos_state_dict as _mk
    return _mk(seed)


def backbone(sd, mel: torch.Tensor, taps=None) -> torch.Tensor:
    """mel (B, 100, T) -> (B, T, 512)."""
    x = F.conv1d(mel, sd["backbone.embed.weight"], sd["backbone.embed.bias"], padding=3)
    x = F.layer_norm(x.transpose(1, 2), (DIM,), sd["backbone.norm.weight"], sd["backbone.norm.bias"], 1e-6)
    x = x.transpose(1, 2)
    if taps is not None:
        taps["embed"] = x.transpose(1, 2)
    for i in range(LAYERS):
        p = f"backbone.convnext.{i}"
        r = x
        x = F.conv1d(x, sd[f"{p}.dwconv.weight"], sd[f"{p}.dwconv.bias"], padding=3, groups=DIM)
        x = F.layer_norm(x.transpose(1, 2), (DIM,), sd[f"{p}.norm.weight"], sd[f"{p}.norm.bias"], 1e-6)
        x = F.gelu(F.linear(x, sd[f"{p}.pwconv1.weight"], sd[f"{p}.pwconv1.bias"]))
        x = F.linear(x, sd[f"{p}.pwconv2.weight"], sd[f"{p}.pwconv2.bias"]) * sd[f"{p}.gamma"]
        x = r + x.transpose(1, 2)
        if taps is not None:
            taps[f"block{i}"] = x.transpose(1, 2)
    return F.layer_norm(x.transpose(1, 2), (DIM,), sd["backbone.final_layer_norm.weight"],
                        sd["backbone.final_layer_norm.bias"], 1e-6)


def head(sd, x: torch.Tensor, taps=None) -> torch.Tensor:
    """(B, T, 512) -> (B, 256 (T-1)) : ISTFTHead with padding="center"."""
    y = F.linear(x, sd["head.out.weight"], sd["head.out.bias"]).transpose(1, 2)
    mag, p = y.chunk(2, dim=1)
    mag = torch.clip(torch.exp(mag), max=1e2)
    S = mag * (torch.cos(p) + 1j * torch.sin(p))
    if taps is not None:
        taps["spec"] = S
    return torch.istft(S, N_FFT, HOP, N_FFT, torch.hann_window(N_FFT), center=True)


def decode(sd, mel: torch.Tensor, taps=None) -> torch.Tensor:
    """Vocos.decode: mel (B, 100, T) -> waveform (B, 256 (T-1))."""
    return head(sd, backbone(sd, mel, taps), taps)

"""Host-side mirror of the reference inference entry points (zipvoice/bin/infer_zipvoice.py):

* ``generate_sentence``  -- same arguments, pre/post scaling and metrics as :276-403, but waveforms /
  token ids are passed in directly (tokenizers and audio file I/O are out of scope);
* ``generate_sentence_stereo`` -- bin/infer_zipvoice_dialog.py:350-515 for ZipVoiceDialogStereo;
* ``generate_batch``     -- what the reference lacks: a whole list of utterances as ONE padded batch
  (GPU mel extraction of all prompts, one ``sample()``, one batched vocoder call), optionally sharded
  data-parallel over the ranks of torchrun and gathered to rank 0.
Everything numerical runs in the CUDA library; this file only moves tensors.
"""
import time
from typing import Dict, List, Optional, Sequence

import torch

from . import parallel


def _rms_normalise(wav: torch.Tensor, target_rms: float):
    """infer_zipvoice.py:340-342: quiet prompts are raised to target_rms; returns (wav, prompt_rms)."""
    rms = torch.sqrt(torch.mean(torch.square(wav)))
    if rms < target_rms:
        wav = wav * target_rms / rms
    return wav, rms


@torch.inference_mode()
def generate_sentence(prompt_tokens: List[int], prompt_wav: torch.Tensor, tokens: List[int], model, vocoder,
                      feature_extractor, num_step: int = 16, guidance_scale: float = 1.0, speed: float = 1.0,
                      t_shift: float = 0.5, target_rms: float = 0.1, feat_scale: float = 0.1,
                      feat_bias: float = 0.0, sampling_rate: int = 24000, x0: Optional[torch.Tensor] = None):
    """-> (wav (1, N) on the device, metrics dict with the reference's keys)."""
    dev = model.device
    prompt_wav = prompt_wav.to(dev).reshape(1, -1).float()
    prompt_wav, prompt_rms = _rms_normalise(prompt_wav, target_rms)
    feats = feature_extractor.extract(prompt_wav, sampling_rate=sampling_rate).to(dev)
    feats = (feats.unsqueeze(0) + feat_bias) * feat_scale                       # :349
    feats_lens = torch.tensor([feats.size(1)], device=dev)
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    pred, pred_lens, _, _ = model.sample(tokens=[tokens], prompt_tokens=[prompt_tokens], prompt_features=feats,
                                         prompt_features_lens=feats_lens, speed=speed, t_shift=t_shift,
                                         duration="predict", num_step=num_step, guidance_scale=guidance_scale,
                                         **({"x0": x0} if x0 is not None else {}))
    pred = pred.permute(0, 2, 1) / feat_scale - feat_bias                       # :374
    torch.cuda.synchronize(dev)
    t1 = time.perf_counter()
    wav = vocoder.decode(pred).clamp(-1, 1)                                     # :378
    torch.cuda.synchronize(dev)
    t2 = time.perf_counter()
    if prompt_rms < target_rms:
        wav = wav * prompt_rms / target_rms                                     # :399-400
    sec = wav.shape[-1] / sampling_rate
    metrics = {"t": t2 - t0, "t_no_vocoder": t1 - t0, "t_vocoder": t2 - t1, "wav_seconds": sec,
               "rtf": (t2 - t0) / sec, "rtf_no_vocoder": (t1 - t0) / sec, "rtf_vocoder": (t2 - t1) / sec}
    return wav, metrics


@torch.inference_mode()
def generate_sentence_stereo(prompt_tokens: List[int], prompt_wav: torch.Tensor, tokens: List[int], model, vocoder,
                             feature_extractor, num_step: int = 16, guidance_scale: float = 1.0, speed: float = 1.0,
                             t_shift: float = 0.5, target_rms: float = 0.1, feat_scale: float = 0.1,
                             sampling_rate: int = 24000, x0: Optional[torch.Tensor] = None):
    """bin/infer_zipvoice_dialog.py:350-515 (generate_sentence_stereo) for a two-channel prompt waveform
    (2, N) already merged / resampled by the caller: RMS over both channels, VocosFbank(num_channels=2)
    features (T, 200), ZipVoiceDialogStereo.sample, one vocoder pass per channel (:483-490).
    -> (wav (2, N') on the device, metrics dict with the reference's keys)."""
    dev = model.device
    prompt_wav = prompt_wav.to(dev).float()
    assert prompt_wav.dim() == 2 and prompt_wav.size(0) == 2, \
        "Merged prompt wav must be stereo for stereo dialogue generation"            # :425-428
    prompt_wav, prompt_rms = _rms_normalise(prompt_wav, target_rms)                   # :445-447
    feats = feature_extractor.extract(prompt_wav, sampling_rate=sampling_rate).to(dev)
    feats = feats.unsqueeze(0) * feat_scale                                           # :454
    feats_lens = torch.tensor([feats.size(1)], device=dev)
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    pred, pred_lens, _, _ = model.sample(tokens=[tokens], prompt_tokens=[prompt_tokens], prompt_features=feats,
                                         prompt_features_lens=feats_lens, speed=speed, t_shift=t_shift,
                                         duration="predict", num_step=num_step, guidance_scale=guidance_scale,
                                         **({"x0": x0} if x0 is not None else {}))
    pred = pred.permute(0, 2, 1) / feat_scale                                         # :479 (B, 2F, T)
    torch.cuda.synchronize(dev)
    t1 = time.perf_counter()
    fd = pred.size(1) // 2
    # both channels in one batched vocoder call (the reference makes two calls, :484-489)
    wav = vocoder.decode(torch.cat([pred[:, :fd], pred[:, fd:2 * fd]], dim=0).contiguous()).clamp(-1, 1)
    torch.cuda.synchronize(dev)
    t2 = time.perf_counter()
    if prompt_rms < target_rms:
        wav = wav * prompt_rms / target_rms                                           # :511-512
    sec = wav.shape[-1] / sampling_rate
    metrics = {"t": t2 - t0, "t_no_vocoder": t1 - t0, "t_vocoder": t2 - t1, "wav_seconds": sec,
               "rtf": (t2 - t0) / sec, "rtf_no_vocoder": (t1 - t0) / sec, "rtf_vocoder": (t2 - t1) / sec}
    return wav, metrics


@torch.inference_mode()
def generate_batch(prompt_tokens: Sequence[List[int]], prompt_wavs: Sequence[torch.Tensor],
                   tokens: Sequence[List[int]], model, vocoder, feature_extractor, num_step: int = 16,
                   guidance_scale: float = 1.0, speed: float = 1.0, t_shift: float = 0.5, target_rms: float = 0.1,
                   feat_scale: float = 0.1, feat_bias: float = 0.0, distributed: bool = False) -> Optional[list]:
    """All utterances of this rank as one padded batch.  With ``distributed`` the list is first sharded by
    (prompt length + text length) over the torchrun ranks and the waveforms are gathered to rank 0, which
    returns them in the original order (other ranks return None)."""
    dev = model.device
    n = len(tokens)
    rank, world = 0, 1
    if distributed and torch.distributed.is_initialized():
        rank, world = torch.distributed.get_rank(), torch.distributed.get_world_size()
    cost = [int(prompt_wavs[i].numel()) // 256 * (1 + len(tokens[i]) // max(1, len(prompt_tokens[i]))) for i in range(n)]
    shards = [parallel.shard_indices(cost, world, r) for r in range(world)]
    mine = shards[rank]
    wav_lens = torch.tensor([int(prompt_wavs[i].numel()) for i in mine], dtype=torch.int64)
    batch = torch.zeros(len(mine), int(wav_lens.max()), device=dev)
    rms = []
    for row, i in enumerate(mine):
        w, r = _rms_normalise(prompt_wavs[i].to(dev).reshape(-1).float(), target_rms)
        batch[row, : w.numel()] = w
        rms.append(r)
    feats = feature_extractor.extract_batch(batch, wav_lens.to(dev))           # (B, Tp, 100), zeros past each prompt
    feats_lens = (wav_lens + 128) // 256
    keep = torch.arange(feats.size(1), device=dev)[None, :] < feats_lens.to(dev)[:, None]
    feats = ((feats + feat_bias) * feat_scale) * keep.unsqueeze(-1)
    pred, pred_lens, _, _ = model.sample(tokens=[tokens[i] for i in mine], prompt_tokens=[prompt_tokens[i] for i in mine],
                                         prompt_features=feats, prompt_features_lens=feats_lens.to(dev), speed=speed,
                                         t_shift=t_shift, duration="predict", num_step=num_step,
                                         guidance_scale=guidance_scale)
    mel = pred.permute(0, 2, 1) / feat_scale - feat_bias
    wav = vocoder.decode(mel, pred_lens).clamp(-1, 1)
    scale = torch.stack([torch.clamp(r / target_rms, max=1.0) for r in rms]).to(dev)
    wav = wav * scale[:, None]
    wlens = (pred_lens - 1).clamp(min=0) * 256
    if world == 1:
        return parallel.scatter_back([(wav, wlens)], shards, n)
    parts = parallel.gather_ragged(wav, wlens, dst=0)
    if rank != 0:
        return None
    return parallel.scatter_back(parts, shards, n)

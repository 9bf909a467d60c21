"""Host-side mirror of the reference inference entry points (zipvoice/bin/infer_zipvoice.py):

* ``generate_sentence``  -- same arguments, pre/post scaling and metrics as :276-403, but waveforms /
  token ids are passed in directly (tokenizers and audio file I/O are out of scope);
* ``generate_sentence_stereo`` -- bin/infer_zipvoice_dialog.py:350-515 for ZipVoiceDialogStereo;
* ``generate_batch``     -- what the reference lacks: a whole list of utterances as ONE padded batch
  (GPU mel extraction of all prompts, one ``sample()``, one batched vocoder call), optionally sharded
  data-parallel over the ranks of torchrun and gathered to rank 0.
* ``generate_list``      -- :406-466 on top of ``generate_batch``: the tab-separated test list is sorted into
  length buckets, every bucket is one padded batch, prompts are resampled / normalised / mel-extracted on the
  GPU, and the 16-bit wav files are written by a background thread while the next bucket runs.
Everything numerical runs in the CUDA library; this file only moves tensors.
"""
import os
import time
import wave
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np
import torch

from . import parallel


def _rms_normalise(wav: torch.Tensor, target_rms: float, rms: Optional[torch.Tensor] = None):
    """infer_zipvoice.py:340-342: quiet prompts are raised to target_rms; returns (wav, prompt_rms).  ``rms`` may be
    handed in when it was measured on another view of the prompt (the un-mixed channels, generate_list).  A silent
    prompt (rms == 0) is left as it is: the reference divides by zero there and feeds NaNs to the model."""
    if rms is None:
        rms = torch.sqrt(torch.mean(torch.square(wav)))
    if 0 < rms < target_rms:
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
                   feat_scale: float = 0.1, feat_bias: float = 0.0, distributed: bool = False,
                   prompt_rms: Optional[Sequence] = None, x0: Optional[torch.Tensor] = None) -> Optional[list]:
    """All utterances of this rank as one padded batch.  With ``distributed`` the list is first sharded by
    (prompt length + text length) over the torchrun ranks and the waveforms are gathered to rank 0, which
    returns them in the original order (other ranks return None).  ``prompt_rms`` (one value per utterance)
    replaces the RMS of the mono prompt, for callers that measured it on the un-mixed channels as the reference does;
    ``x0`` (B, T, F) injects the start noise (single-rank only; rows in the order of the arguments).

    Output is tolerance-level, not bit-level, equal to per-utterance ``generate_sentence`` calls: Zipformer's
    down- and up-sampling mix the last valid frames of an utterance with the padded frames behind it, and the
    reference draws different noise per call."""
    dev = model.device
    n = len(tokens)
    rank, world = 0, 1
    if distributed and torch.distributed.is_initialized():
        rank, world = torch.distributed.get_rank(), torch.distributed.get_world_size()
    cost = [int(prompt_wavs[i].numel()) // 256 * (1 + len(tokens[i]) // max(1, len(prompt_tokens[i]))) for i in range(n)]
    shards = [parallel.shard_indices(cost, world, r) for r in range(world)]
    mine = shards[rank]
    if not mine:  # more ranks than utterances: this rank only takes part in the gather
        parts = parallel.gather_ragged(torch.zeros(0, 0, device=dev), torch.zeros(0, dtype=torch.int64, device=dev))
        return parallel.scatter_back(parts, shards, n) if (world == 1 or rank == 0) else None
    assert x0 is None or world == 1, "x0 is given in argument order: single-rank calls only"
    wav_lens = torch.tensor([int(prompt_wavs[i].numel()) for i in mine], dtype=torch.int64)
    batch = torch.zeros(len(mine), int(wav_lens.max()), device=dev)
    rms = []
    for row, i in enumerate(mine):
        given = None if prompt_rms is None else torch.as_tensor(prompt_rms[i], dtype=torch.float32, device=dev)
        w, r = _rms_normalise(prompt_wavs[i].to(dev).reshape(-1).float(), target_rms, given)
        batch[row, : w.numel()] = w
        rms.append(r)
    feats = feature_extractor.extract_batch(batch, wav_lens.to(dev))           # (B, Tp, 100), zeros past each prompt
    feats_lens = (wav_lens + 128) // 256
    keep = torch.arange(feats.size(1), device=dev)[None, :] < feats_lens.to(dev)[:, None]
    feats = ((feats + feat_bias) * feat_scale) * keep.unsqueeze(-1)
    pred, pred_lens, _, _ = model.sample(tokens=[tokens[i] for i in mine], prompt_tokens=[prompt_tokens[i] for i in mine],
                                         prompt_features=feats, prompt_features_lens=feats_lens.to(dev), speed=speed,
                                         t_shift=t_shift, duration="predict", num_step=num_step,
                                         guidance_scale=guidance_scale,
                                         **({"x0": x0[mine]} if x0 is not None else {}))  # batch rows are in shard (length-sorted) order
    mel = pred.permute(0, 2, 1) / feat_scale - feat_bias
    wav = vocoder.decode(mel, pred_lens).clamp(-1, 1)
    # infer_zipvoice.py:399-400: only prompts that were raised are lowered again (a silent prompt was not raised)
    scale = torch.stack([torch.where(r > 0, torch.clamp(r / target_rms, max=1.0), torch.ones_like(r))
                         for r in rms]).to(dev)
    wav = wav * scale[:, None]
    wlens = (pred_lens - 1).clamp(min=0) * 256
    if world == 1:
        return parallel.scatter_back([(wav, wlens)], shards, n)
    parts = parallel.gather_ragged(wav, wlens, dst=0)
    if rank != 0:
        return None
    return parallel.scatter_back(parts, shards, n)


def read_wav(path: str):
    """PCM wav file -> ((channels, samples) float32 in [-1, 1), sampling rate); 16- and 32-bit integer samples.
    Stand-in for torchaudio.load (infer_zipvoice.py:332) on machines without an audio backend."""
    with wave.open(path, "rb") as f:
        sr, ch, width, n = f.getframerate(), f.getnchannels(), f.getsampwidth(), f.getnframes()
        raw = f.readframes(n)
    if width == 2:
        x = np.frombuffer(raw, dtype="<i2").astype(np.float32) / 32768.0
    elif width == 4:
        x = np.frombuffer(raw, dtype="<i4").astype(np.float32) / 2147483648.0
    else:
        raise ValueError(f"{path}: unsupported sample width {width}")
    return torch.from_numpy(x.reshape(-1, ch).T.copy()), sr


def write_wav(path: str, wav: np.ndarray, sampling_rate: int):
    """float waveform in [-1, 1], (N,) mono or (C, N) -> 16-bit PCM with interleaved channels (what torchaudio.save
    writes for the reference, infer_zipvoice.py:401, infer_zipvoice_dialog.py:346,513)."""
    w = np.asarray(wav, dtype=np.float32)
    w = w.reshape(1, -1) if w.ndim == 1 else w
    pcm = np.clip(np.round(w.T.reshape(-1) * 32768.0), -32768, 32767).astype("<i2")
    with wave.open(path, "wb") as f:
        f.setnchannels(int(w.shape[0]))
        f.setsampwidth(2)
        f.setframerate(int(sampling_rate))
        f.writeframes(pcm.tobytes())


def merge_dialog_prompts(wavs: Sequence[torch.Tensor], stereo: bool,
                         silence: Optional[torch.Tensor] = None) -> torch.Tensor:
    """The prompt assembly of the dialog CLI on waveforms already at the model's sampling rate, each (C, N).
    Mono model (infer_zipvoice_dialog.py:268-279): every file is down-mixed to one channel, two files are joined
    along time -> (1, N).  Stereo model (:424-441): one file must have two channels; two two-channel files are joined
    along time; two one-channel files go to channel 0 / channel 1 one after the other on top of the ``silence``
    template (2, >= N1 + N2) -> (2, N)."""
    wavs = [w.reshape(1, -1) if w.dim() == 1 else w for w in wavs]
    assert len(wavs) in (1, 2)
    if not stereo:
        return torch.cat([w.mean(0, keepdim=True) if w.size(0) != 1 else w for w in wavs], dim=1)
    if len(wavs) == 1:
        assert wavs[0].size(0) == 2, "Merged prompt wav must be stereo for stereo dialogue generation"
        return wavs[0]
    if wavs[0].size(0) == 2:
        return torch.cat(list(wavs), dim=1)
    assert wavs[0].size(0) == 1 and silence is not None, "two mono prompts need the silence template (silence_wav)"
    n1, n2 = wavs[0].size(1), wavs[1].size(1)
    out = silence[:, : n1 + n2].clone()
    out[0, :n1] = wavs[0][0]
    out[1, n1:] = wavs[1][0]
    return out


def length_buckets(cost: Sequence[int], max_batch_cost: int) -> List[List[int]]:
    """Indices sorted by cost and cut into batches whose padded size (longest x count) stays within the budget."""
    order = sorted(range(len(cost)), key=lambda i: cost[i])
    out, cur = [], []
    for i in order:
        if cur and cost[i] * (len(cur) + 1) > max_batch_cost:   # ascending order: cost[i] is the batch's longest
            out.append(cur)
            cur = []
        cur.append(i)
    if cur:
        out.append(cur)
    return out


@torch.inference_mode()
def generate_list(res_dir: str, test_list: str, model, vocoder, tokenizer, feature_extractor, num_step: int = 16,
                  guidance_scale: float = 1.0, speed: float = 1.0, t_shift: float = 0.5, target_rms: float = 0.1,
                  feat_scale: float = 0.1, feat_bias: float = 0.0, sampling_rate: int = 24000,
                  max_batch_frames: int = 64 * 1300, distributed: bool = False,
                  load_wav: Callable = read_wav) -> Dict[str, float]:
    """infer_zipvoice.py:406-466.  ``test_list`` lines: wav_name <tab> prompt_text <tab> prompt_wav <tab> text;
    ``tokenizer`` needs ``texts_to_token_ids`` like the reference's tokenizers.  Writes ``res_dir/<wav_name>.wav``.

    Differences from the reference loop: utterances run as padded batches of similar length (``max_batch_frames``
    bounds batch x frames) with their own noise, so the output equals the per-utterance loop to tolerance, not bit
    for bit (see ``generate_batch``); prompts of one sampling rate are resampled together on the GPU (all channels,
    one row each); the prompt RMS is measured on the resampled un-mixed channels exactly as :340-342 does, and the
    channel mean that VocosFbank applies (feature.py:84-86) is taken right after it; files are written behind the
    compute.  With ``distributed`` the list is split over the torchrun ranks by estimated cost and every rank
    writes its own files (no collective on the data path).  Returns the totals the reference logs."""
    from .audio import ResampleB200
    dev = model.device
    with open(test_list, "r") as fr:
        items = [ln.rstrip("\n").split("\t") for ln in fr if ln.strip()]
    for it in items:
        if len(it) != 4:
            raise ValueError(f"malformed test-list line (need 4 tab-separated fields): {it}")
    os.makedirs(res_dir, exist_ok=True)
    rank, world = 0, 1
    if distributed and torch.distributed.is_initialized():
        rank, world = torch.distributed.get_rank(), torch.distributed.get_world_size()
    tokens = tokenizer.texts_to_token_ids([it[3] for it in items])
    prompt_tokens = tokenizer.texts_to_token_ids([it[1] for it in items])
    # rank split on a cost known before any audio is read: characters to synthesise
    text_cost = [len(t) + len(p) for t, p in zip(tokens, prompt_tokens)]
    mine = parallel.shard_indices(text_cost, world, rank)
    # ---- prompts: load, resample per source rate as one padded GPU batch (one row per channel), RMS, down-mix
    loaded = {i: load_wav(items[i][2]) for i in mine}
    prompts: Dict[int, torch.Tensor] = {}
    prompt_rms: Dict[int, torch.Tensor] = {}
    by_rate: Dict[int, List[int]] = {}
    for i, (w, sr) in loaded.items():
        by_rate.setdefault(int(sr), []).append(i)
    for sr, idx in by_rate.items():
        chans = {i: loaded[i][0].float().reshape(-1, loaded[i][0].shape[-1]) for i in idx}   # (C_i, N_i)
        if sr != sampling_rate:
            rows = [(i, c) for i in idx for c in range(chans[i].shape[0])]
            lens = torch.tensor([chans[i].shape[1] for i, _ in rows], dtype=torch.int64)
            batch = torch.zeros(len(rows), int(lens.max()))
            for r, (i, c) in enumerate(rows):
                batch[r, : chans[i].shape[1]] = chans[i][c]
            res = ResampleB200(sr, sampling_rate, engine=feature_extractor.engine)(batch.to(dev), lens.to(dev))
            out_lens = (lens * sampling_rate + sr - 1) // sr
            got: Dict[int, List[torch.Tensor]] = {}
            for r, (i, c) in enumerate(rows):
                got.setdefault(i, []).append(res[r, : int(out_lens[r])])
            chans = {i: torch.stack(v) for i, v in got.items()}
        for i in idx:
            w = chans[i].to(dev)
            prompt_rms[i] = torch.sqrt(torch.mean(torch.square(w)))     # :340, on every channel
            prompts[i] = w.mean(dim=0)                                  # feature.py:84-86 (num_channels = 1)
    # ---- buckets of similar total length (prompt frames scaled by the text / prompt-text ratio)
    frames = {i: max(1, prompts[i].numel() // 256) for i in mine}
    cost = [int(frames[i] * (1 + len(tokens[i]) / max(1, len(prompt_tokens[i])) / max(speed, 1e-3))) for i in mine]
    buckets = [[mine[j] for j in b] for b in length_buckets(cost, max_batch_frames)]
    total_t, total_sec, n_done = 0.0, 0.0, 0
    pending = []
    with ThreadPoolExecutor(max_workers=2) as pool:
        for b in buckets:
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            wavs = generate_batch([prompt_tokens[i] for i in b], [prompts[i] for i in b], [tokens[i] for i in b],
                                  model, vocoder, feature_extractor, num_step=num_step,
                                  guidance_scale=guidance_scale, speed=speed, t_shift=t_shift, target_rms=target_rms,
                                  feat_scale=feat_scale, feat_bias=feat_bias, prompt_rms=[prompt_rms[i] for i in b])
            host = [w.cpu().numpy() for w in wavs]          # device -> host read of the bucket's waveforms
            total_t += time.perf_counter() - t0
            for i, w in zip(b, host):
                total_sec += w.shape[-1] / sampling_rate
                pending.append(pool.submit(write_wav, os.path.join(res_dir, f"{items[i][0]}.wav"), w, sampling_rate))
            n_done += len(b)
        for f in pending:
            f.result()
    return {"t": total_t, "wav_seconds": total_sec, "rtf": total_t / max(total_sec, 1e-9), "utterances": n_done,
            "batches": len(buckets)}


@torch.inference_mode()
def generate_list_dialog(model_name: str, res_dir: str, test_list: str, model, vocoder, tokenizer, feature_extractor,
                         num_step: int = 16, guidance_scale: float = 1.5, speed: float = 1.0, t_shift: float = 0.5,
                         target_rms: float = 0.1, feat_scale: float = 0.1, sampling_rate: int = 24000,
                         silence_wav: Optional[str] = None, max_batch_frames: int = 64 * 1300,
                         load_wav: Callable = read_wav) -> Dict[str, float]:
    """bin/infer_zipvoice_dialog.py:518-611.  ``test_list`` lines are either
        wav_name <tab> prompt_text <tab> prompt_wav <tab> text                                   (merged prompt), or
        wav_name <tab> prompt_text_1 <tab> prompt_text_2 <tab> prompt_wav_1 <tab> prompt_wav_2 <tab> text
    (the prompt text becomes "[S1]...[S2]...", :546); ``text`` must start with "[S1]" (:552).
    ``model_name`` "zipvoice_dialog": prompts are resampled per file, down-mixed and joined along time
    (``merge_dialog_prompts``), the RMS is that of the joined mono prompt (:280-282), and the utterances run as padded
    length buckets through ``generate_batch`` (tolerance-level equal to the reference's per-utterance loop, see there).
    "zipvoice_dialog_stereo": two-channel prompts and outputs through ``generate_sentence_stereo`` per utterance.
    Writes ``res_dir/<wav_name>.wav`` (two channels for the stereo model) and returns the totals the reference logs."""
    from .audio import ResampleB200
    assert model_name in ("zipvoice_dialog", "zipvoice_dialog_stereo"), model_name
    stereo = model_name == "zipvoice_dialog_stereo"
    dev = model.device
    names, ptexts, pwavs, texts = [], [], [], []
    with open(test_list, "r") as fr:
        for line in fr:
            if not line.strip():
                continue
            items = line.strip().split("\t")
            if len(items) == 6:
                wav_name, pt1, pt2, pw1, pw2, text = items
                ptext, pw = f"[S1]{pt1}[S2]{pt2}", [pw1, pw2]
            elif len(items) == 4:
                wav_name, ptext, pw1, text = items
                pw = [pw1]
            else:
                raise ValueError(f"Invalid line: {line}")
            assert text.startswith("[S1]")
            names.append(wav_name), ptexts.append(ptext), pwavs.append(pw), texts.append(text)
    os.makedirs(res_dir, exist_ok=True)
    tokens = tokenizer.texts_to_token_ids(texts)
    prompt_tokens = tokenizer.texts_to_token_ids(ptexts)
    silence = None
    if stereo and silence_wav is not None:
        silence, ssr = load_wav(silence_wav)
        assert ssr == sampling_rate                                              # :436
    resamplers: Dict[int, ResampleB200] = {}

    def load(path):  # :268-277 / :416-422: load, resample the file as it is (all channels) to the model's rate
        w, sr = load_wav(path)
        w = w.float().reshape(-1, w.shape[-1]).to(dev)
        if int(sr) != sampling_rate:
            if int(sr) not in resamplers:
                resamplers[int(sr)] = ResampleB200(int(sr), sampling_rate, engine=feature_extractor.engine)
            w = resamplers[int(sr)](w)
        return w

    prompts = [merge_dialog_prompts([load(pth) for pth in pw], stereo, None if silence is None else silence.to(dev))
               for pw in pwavs]
    total_t, total_sec = 0.0, 0.0
    pending = []
    with ThreadPoolExecutor(max_workers=2) as pool:
        if stereo:
            for i, name in enumerate(names):
                wav, m = generate_sentence_stereo(prompt_tokens[i], prompts[i], tokens[i], model, vocoder,
                                                  feature_extractor, num_step=num_step, guidance_scale=guidance_scale,
                                                  speed=speed, t_shift=t_shift, target_rms=target_rms,
                                                  feat_scale=feat_scale, sampling_rate=sampling_rate)
                total_t += m["t"]
                total_sec += m["wav_seconds"]
                pending.append(pool.submit(write_wav, os.path.join(res_dir, f"{name}.wav"), wav.cpu().numpy(),
                                           sampling_rate))
            n_batches = len(names)
        else:
            frames = [max(1, p.shape[-1] // 256) for p in prompts]
            cost = [int(frames[i] * (1 + len(tokens[i]) / max(1, len(prompt_tokens[i])) / max(speed, 1e-3)))
                    for i in range(len(names))]
            buckets = length_buckets(cost, max_batch_frames)
            for b in buckets:
                torch.cuda.synchronize(dev)
                t0 = time.perf_counter()
                wavs = generate_batch([prompt_tokens[i] for i in b], [prompts[i][0] for i in b], [tokens[i] for i in b],
                                      model, vocoder, feature_extractor, num_step=num_step,
                                      guidance_scale=guidance_scale, speed=speed, t_shift=t_shift,
                                      target_rms=target_rms, feat_scale=feat_scale, feat_bias=0.0)
                host = [w.cpu().numpy() for w in wavs]
                total_t += time.perf_counter() - t0
                for i, w in zip(b, host):
                    total_sec += w.shape[-1] / sampling_rate
                    pending.append(pool.submit(write_wav, os.path.join(res_dir, f"{names[i]}.wav"), w, sampling_rate))
            n_batches = len(buckets)
        for f in pending:
            f.result()
    return {"t": total_t, "wav_seconds": total_sec, "rtf": total_t / max(total_sec, 1e-9), "utterances": len(names),
            "batches": n_batches}

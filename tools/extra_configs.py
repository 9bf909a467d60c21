#!/usr/bin/env python
"""Measurements of BASELINE.json configs 1, 3, 4, 5 on one GPU (config 2 is bench.py itself).

    python tools/extra_configs.py [--out gpurun_out/extra_configs.json] [--skip-cpu]

config 1: B=1, 75-token text, 3 s prompt, 16-step CFG sample -- this repo's CUDA path and the CPU oracle port.
config 3: ZipVoice-Distill, 32 utterances per GPU (256 over 8 GPUs), 4 NFE, guidance 3.0 embedded, T=1219.
config 4: ZipVoice-Dialog, B=1, 10 s prompt + 60 s target (T=6563), 16 NFE, CFG 1.5.
config 5: mel front-end and vocoder on clips of 1/5/10/30/60 s x batch {1, 8, 64, 512} (B x N capped by memory).
All timings: CUDA events, 2 warm-ups, 3 repetitions, mean.  Synthetic seeded inputs, random-init weights.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from zipvoice_b200 import synth  # noqa: E402
from zipvoice_b200.model import ZipVoiceB200  # noqa: E402

FRAME_S = 256.0 / 24000.0
dev = torch.device("cuda:0")


def timed(fn, reps=3, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def sample_case(kind, B, n_prompt_tok, n_text_tok, prompt_frames, target_frames, num_step, guidance, t_shift):
    cfg = synth.BASE_MODEL_CFG
    sd = synth.make_state_dict(cfg, seed=0, distill=(kind == "distill"), dialog=(kind == "dialog"))
    model = ZipVoiceB200(cfg, sd, synth.VOCAB_SIZE, synth.PAD_ID, kind=kind, device=dev)
    tokens, prompt_tokens = synth.make_tokens(B, n_prompt_tok, n_text_tok, seed=1234, dialog=(kind == "dialog"))
    pf, pfl = synth.make_prompt_features(B, prompt_frames, seed=4321)
    T = prompt_frames + target_frames
    x0 = synth.make_noise(B, T, seed=666).to(dev)
    flens = torch.full((B,), target_frames, dtype=torch.int64)
    text_cond, pad_mask = model.forward_text_inference_gt_duration(tokens, flens.to(dev), prompt_tokens, pfl.to(dev))
    speech = torch.zeros(B, T, 100, device=dev)
    speech[:, :prompt_frames] = pf.to(dev)
    eng = model.engine

    def step():
        return eng.sample(x0, text_cond, speech, pad_mask, num_step=num_step, guidance_scale=guidance, t_shift=t_shift)

    ms = timed(step)
    out = step()
    assert torch.isfinite(out).all()
    audio_s = B * target_frames * FRAME_S
    res = {"kind": kind, "B": B, "T": T, "nfe": num_step, "guidance": guidance, "ms_per_sample_call": ms,
           "ms_per_nfe": ms / num_step, "audio_s_per_s": audio_s / (ms * 1e-3)}
    eng.close()
    del model
    torch.cuda.empty_cache()
    return res


def cpu_config1(threads):
    """CPU oracle port of the reference on config 1 (B=1, CFG batch 2): 2 timed NFEs -> 16-NFE sample time."""
    from oracle import zipvoice_oracle as O
    from tests.cases import case_inputs
    torch.set_num_threads(threads)
    cfg = synth.BASE_MODEL_CFG
    sd = synth.make_state_dict(cfg, seed=0)
    T = 281 + 469
    x0, text, speech, mask = case_inputs(1, T, 281, seed=21, ragged=False)
    ts = []
    with torch.inference_mode():
        for _ in range(3):
            t0 = time.perf_counter()
            O.cfg_velocity(sd, cfg, torch.tensor(0.3), x0, text, speech, mask, 1.0)
            ts.append(time.perf_counter() - t0)
    per_nfe = min(ts[1:])
    return {"threads": threads, "s_per_nfe": per_nfe, "audio_s_per_s": 469 * FRAME_S / (16 * per_nfe)}


def audio_sweep():
    from zipvoice_b200.audio import audio_engine
    eng = audio_engine(synth.make_vocos_state_dict(0), device=dev)
    rows = []
    for secs in (1, 5, 10, 30, 60):
        n = secs * 24000
        for B in (1, 8, 64, 512):
            if B * n > 64 * 60 * 24000:  # cap: 64 minutes of audio per call (vocoder activations ~ 14 KB per frame)
                continue
            wav = synth.make_audio(B, n, seed=7).to(dev)
            ms_f = timed(lambda: eng.fbank(wav))
            mel = eng.fbank(wav)  # (B, T, 100)
            melT = mel.permute(0, 2, 1).contiguous()
            ms_v = timed(lambda: eng.vocos_decode(melT))
            frames = mel.shape[1]
            rows.append({"clip_s": secs, "B": B, "frames": frames, "fbank_ms": ms_f, "vocoder_ms": ms_v,
                         "fbank_audio_s_per_s": B * secs / (ms_f * 1e-3), "vocoder_audio_s_per_s": B * secs / (ms_v * 1e-3),
                         "fbank_GBps_algorithmic": B * (n * 4 + frames * 400) / (ms_f * 1e-3) / 1e9,
                         "vocoder_TFLOPs_gemm": B * frames * 27.0e6 / (ms_v * 1e-3) / 1e12})
            del wav, mel, melT
            torch.cuda.empty_cache()
    eng.close()
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "extra_configs.json"))
    ap.add_argument("--skip-cpu", action="store_true")
    args = ap.parse_args()
    res = {"gpu": torch.cuda.get_device_name(0)}
    res["config1_gpu"] = sample_case("zipvoice", 1, 45, 75, 281, 469, 16, 1.0, 0.5)
    res["config3_distill_per_gpu"] = sample_case("distill", 32, 45, 150, 281, 938, 4, 3.0, 0.5)
    res["config3_distill_b256_one_gpu"] = sample_case("distill", 256, 45, 150, 281, 938, 4, 3.0, 0.5)
    res["config4_dialog"] = sample_case("dialog", 1, 150, 900, 938, 5625, 16, 1.5, 0.5)
    res["config5_audio"] = audio_sweep()
    if not args.skip_cpu:
        res["config1_cpu_oracle_1thread"] = cpu_config1(1)
        res["config1_cpu_oracle_allcores"] = cpu_config1(os.cpu_count() or 1)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(res, open(args.out, "w"), indent=1)
    for k, v in res.items():
        if k == "config5_audio":
            for r in v:
                print("config5", {a: (round(b, 3) if isinstance(b, float) else b) for a, b in r.items()})
        else:
            print(k, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items()} if isinstance(v, dict) else v)


if __name__ == "__main__":
    main()

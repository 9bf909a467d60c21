#!/usr/bin/env python
"""Benchmark of the ZipVoice sampling hot path on B200 (contract: see DESIGN.md "Measurement").

    python bench.py [--config {1,2,3,4,5}] --gpus N --steps K --warmup W     # this repo's CUDA path
    python bench.py --impl reference [--config k] --gpus N --steps K ...     # CPU reference arm (oracle port)

--config selects BASELINE.json configs[k-1] (default 2 = the headline):
  1  ZipVoice 123M, one ~5 s utterance (B = 1, 3 s prompt + 5 s target, T = 750), 16 NFE, CFG 1.0
  2  ZipVoice 123M, B = 64 per GPU, 3 s prompt + 10 s target (T = 1219), 16 NFE, CFG 1.0 (batch doubled to 128)
  3  ZipVoice-Distill, 32 utterances per GPU (256 over 8 GPUs), 4 NFE, guidance 3.0 embedded, T = 1219
  4  ZipVoice-Dialog, B = 1, 10 s prompt + 60 s target (T = 6563), 16 NFE, CFG 1.5 (N > 1: independent replicas)
  5  VocosFbank mel extraction + Vocos vocoder sweep, 24 kHz clips of 1-60 s x batch 1-512 (capped at 64 min / call)
--ragged (configs 2, 3): targets of 6-10 s (seed 99) instead of 10 s; the world * B utterances are dealt to the ranks
  by length (parallel.shard_indices) and the e2e leg gathers padded mels AND waveforms (parallel.gather_ragged).

One "step" = one full pass of the hot path over one batch (configs 1-4: ``EulerSolver.sample``; config 5: one pass of
the whole sweep).  Metric: generated audio-seconds per second, whole job over all ranks; also ms per NFE.

  value : inputs resident in HBM, timed with CUDA events around K steps (barrier + sync both sides, max over ranks).
  e2e   : the same batch through the public API (``ZipVoiceB200.sample`` -- text encoder, text condition, prompt padding,
          solver, prompt strip -- + the vocoder; config 5: ``fbank`` + ``vocos_decode``) from pinned HOST buffers to HOST
          waveforms, H2D and D2H copies inside the timed region; N > 1 adds the NCCL gather of mels and waveforms.
  roofline / kernels : one extra instrumented step in which the library brackets every launch of its kernels with
          CUDA events on the launching stream (zv_prof_*).
  cpu_baseline / --impl reference : the CPU oracle (torch-fp32 port of the reference, oracle/; kind "port": the
          reference itself is Python and /root/reference does not exist on the GPU box) timed on the host cores on
          a bounded sample of the same workload: ONE utterance of the batch through text encoder, n guided NFEs
          (n <= num_step, scaled to the full num_step; the scale is declared in the line) and the vocoder.
"""
import argparse
import glob
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FRAME_S = 256.0 / 24000.0

CONFIGS = {
    1: dict(kind="zipvoice", B=1, n_prompt_tok=45, n_text_tok=75, prompt_frames=281, target_frames=469, nfe=16,
            guidance=1.0, t_shift=0.5,
            workload="configs[0]: ZipVoice 123M random-init, one ~5 s utterance (B=1, 3 s prompt + 5 s target, T=750, "
                     "100-bin 24 kHz mel), 16 NFE Euler, CFG guidance 1.0 (batch doubled to 2), t_shift 0.5"),
    2: dict(kind="zipvoice", B=64, n_prompt_tok=45, n_text_tok=150, prompt_frames=281, target_frames=938, nfe=16,
            guidance=1.0, t_shift=0.5,
            workload="configs[1]: ZipVoice 123M random-init, B=64 per GPU, 3 s prompt + 10 s target (T=1219, 100-bin "
                     "24 kHz mel), 16 NFE Euler, CFG guidance 1.0 (batch doubled to 128), t_shift 0.5"),
    3: dict(kind="distill", B=32, n_prompt_tok=45, n_text_tok=150, prompt_frames=281, target_frames=938, nfe=4,
            guidance=3.0, t_shift=0.5,
            workload="configs[2]: ZipVoice-Distill 123M random-init, 32 utterances per GPU (256 over 8 GPUs, sharded "
                     "data-parallel), 3 s prompt + 10 s target (T=1219), 4 NFE, guidance 3.0 embedded (no batch doubling)"),
    4: dict(kind="dialog", B=1, n_prompt_tok=150, n_text_tok=900, prompt_frames=938, target_frames=5625, nfe=16,
            guidance=1.5, t_shift=0.5,
            workload="configs[3]: ZipVoice-Dialog 123M random-init, B=1, two-speaker dialogue, 10 s prompt + 60 s target "
                     "(T=6563), 16 NFE Euler, CFG guidance 1.5 (batch doubled to 2); N>1: independent replicas"),
    5: dict(kind="audio", workload="configs[4]: VocosFbank mel extraction + Vocos vocoder sweep on synthetic 24 kHz "
                                   "audio, clips of 1/5/10/30/60 s x batch 1/8/64/512 (B x N capped at 64 minutes of "
                                   "audio per call), every rank runs the whole sweep on its own clips"),
}
AUDIO_SWEEP = [(s, b) for s in (1, 5, 10, 30, 60) for b in (1, 8, 64, 512) if s * b <= 64 * 60]


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d.get("bf16_tflops_sustained",
                    d["bf16_tflops"]), source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=open(self.path, "w"),
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])), mx.append(float(f[2])), pw.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ workload
def target_lengths(c: dict, n_utt: int, ragged: bool):
    """Target (generated) frames of the job's n_utt utterances: 10 s each, or 6-10 s uniformly (seed 99)."""
    import torch
    if not ragged:
        return [c["target_frames"]] * n_utt
    g = torch.Generator().manual_seed(99)
    lo = int(round(c["target_frames"] * 0.6))
    return torch.randint(lo, c["target_frames"] + 1, (n_utt,), generator=g).tolist()


def config_dict(k: int, c: dict, B: int, world: int, ragged: bool) -> dict:
    """The `config` object of the JSON line -- identical for the GPU arm and the reference arm."""
    d = {"workload": c["workload"], "config_index": k, "parallelism": f"dp{world} over utterances"}
    if c["kind"] == "audio":
        d.update({"sweep_clip_s_x_batch": AUDIO_SWEEP, "audio_s_per_pass_per_gpu": sum(s * b for s, b in AUDIO_SWEEP),
                  "l2": "every case but the two smallest moves more than the 126 MB L2 per kernel; no explicit flush"})
        return d
    d.update({"model": c["kind"], "batch_per_gpu": B, "frames": c["prompt_frames"] + c["target_frames"],
              "nfe": c["nfe"], "guidance_scale": c["guidance"], "t_shift": c["t_shift"],
              "weights": "synth.make_state_dict(seed=0), 123M",
              "lengths": "ragged: targets 6-10 s (seed 99), length-balanced shards" if ragged else "equal",
              "l2": ("activations of one NFE (several GB) exceed the 126 MB L2; no explicit flush needed" if B >= 8 else
                     "weights (245 MB bf16) exceed the 126 MB L2 and are re-streamed every NFE; activations stay in L2")})
    return d


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_sample_time(c: dict, nfe_timed: int, threads: int, warm: int = 1):
    """Times the CPU oracle port on ONE utterance of config `c`: text encoder + text condition (once), `nfe_timed`
    guided velocity evaluations of the Euler loop (CFG: batch 1 -> 2), the vocoder (once).
    -> dict(t_text, t_nfe (median seconds per NFE), t_voc, audio_s)."""
    import torch
    from oracle import vocos_oracle as VO
    from oracle import zipvoice_oracle as O
    from zipvoice_b200 import synth
    torch.set_num_threads(threads)
    kind = c["kind"]
    cfg = synth.BASE_MODEL_CFG
    sd = synth.make_state_dict(cfg, seed=0, distill=(kind == "distill"), dialog=(kind == "dialog"))
    vsd = synth.make_vocos_state_dict(0)
    tokens, prompt_tokens = synth.make_tokens(1, c["n_prompt_tok"], c["n_text_tok"], seed=1234, dialog=(kind == "dialog"))
    pf, pfl = synth.make_prompt_features(1, c["prompt_frames"], seed=4321)
    T = c["prompt_frames"] + c["target_frames"]
    x = synth.make_noise(1, T, seed=666)
    dialog = (synth.SPK_A_ID, synth.SPK_B_ID) if kind == "dialog" else None
    ts = O.get_time_steps(0.0, 1.0, c["nfe"], c["t_shift"])
    with torch.inference_mode():
        t0 = time.perf_counter()
        embed, cat_lens = O.forward_text_embed(sd, cfg, [prompt_tokens[0] + tokens[0]], synth.PAD_ID, dialog)
        text, mask = O.forward_text_condition(embed, cat_lens, torch.tensor([T]))
        t_text = time.perf_counter() - t0
        speech = torch.nn.functional.pad(pf, (0, 0, 0, T - pf.size(1)))
        fn = O.distill_velocity if kind == "distill" else O.cfg_velocity
        per = []
        for i in range(warm + nfe_timed):
            step = min(max(i - warm, 0), c["nfe"] - 1)
            t0 = time.perf_counter()
            v = fn(sd, cfg, ts[step], x, text, speech, mask, c["guidance"])
            per.append(time.perf_counter() - t0)
            if i >= warm:
                x = x + v * (ts[step + 1] - ts[step])
        t0 = time.perf_counter()
        VO.decode(vsd, (x[:, c["prompt_frames"]:] / 0.1).permute(0, 2, 1))
        t_voc = time.perf_counter() - t0
    return dict(t_text=t_text, t_nfe=statistics.median(per[warm:]), t_voc=t_voc, audio_s=c["target_frames"] * FRAME_S)


def cpu_audio_time(threads: int, clip_s: int = 10):
    """Config 5 on the CPU: the fbank oracle (torchaudio MelSpectrogram semantics) + the vocoder oracle on one clip."""
    import torch
    from oracle import fbank_oracle as FO
    from oracle import vocos_oracle as VO
    from zipvoice_b200 import synth
    torch.set_num_threads(threads)
    vsd = synth.make_vocos_state_dict(0)
    wav = synth.make_audio(1, clip_s * 24000, seed=7)
    with torch.inference_mode():
        FO.vocos_fbank(wav[0])
        t0 = time.perf_counter()
        mel = FO.vocos_fbank(wav[0])
        VO.decode(vsd, mel.t().unsqueeze(0))
        dt = time.perf_counter() - t0
    return dt, clip_s


def cpu_baseline_obj(k: int, c: dict, threads: int):
    if c["kind"] == "audio":
        dt, clip_s = cpu_audio_time(threads)
        return {"value": clip_s / dt, "unit": "audio-s/s", "cores": threads, "kind": "port",
                "sample": f"one {clip_s} s clip: fbank oracle + vocoder oracle (torch fp32, parity of the vocoder oracle "
                          f"unpinned), 1 warm-up + 1 timed pass"}
    n = 3 if c["target_frames"] < 2000 else 1
    r = cpu_sample_time(c, nfe_timed=n, threads=threads)
    total = r["t_text"] + c["nfe"] * r["t_nfe"] + r["t_voc"]
    one = {}
    if k == 1:  # also at 1 thread, the reference CLI's own default (infer_zipvoice.py:637-638)
        r1 = cpu_sample_time(c, nfe_timed=2, threads=1)
        one = {"value_1_thread": r1["audio_s"] / (r1["t_text"] + c["nfe"] * r1["t_nfe"] + r1["t_voc"]),
               "s_per_nfe_1_thread": r1["t_nfe"]}
    return {"value": r["audio_s"] / total, "unit": "audio-s/s", "cores": threads, "kind": "port", **one,
            "s_per_nfe": r["t_nfe"], "s_text_encoder": r["t_text"], "s_vocoder": r["t_voc"],
            "sample": f"1 utterance of the batch (B=1{' -> CFG batch 2' if c['kind'] != 'distill' else ''}, "
                      f"T={c['prompt_frames'] + c['target_frames']}) through the torch-fp32 CPU oracle port: text encoder + "
                      f"1 warm-up and {n} timed guided NFE (median, scaled x{c['nfe']}) + vocoder oracle; "
                      f"audio-s/s = target seconds / (t_text + {c['nfe']} t_nfe + t_vocoder)"}


def run_reference(args):
    """Reference arm: the CPU oracle port (kind "port") on all host cores, same `config` object as the GPU arm.
    One step = ONE utterance of the batch through text encoder + n NFEs of the Euler loop + vocoder, n <= num_step
    chosen so that K + W steps end within ~4 minutes; the line declares the scale (`sample_scale`)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    threads = os.cpu_count() or 1
    K, W = max(1, args.steps), max(0, args.warmup)
    k, c = args.config, CONFIGS[args.config]
    B = args.batch or c.get("B", 0)
    cfgd = config_dict(k, c, B, world, args.ragged)
    if c["kind"] == "audio":
        ts = []
        for _ in range(min(W, 1) + K):
            dt, clip_s = cpu_audio_time(threads)
            ts.append(dt)
        dt = statistics.mean(ts[min(W, 1):])
        value = clip_s / dt
        sample = f"each step = one {clip_s} s clip through the fbank oracle + the (unpinned) vocoder oracle"
        scale = {"audio_s_timed_per_step": clip_s, "audio_s_per_step_in_config": cfgd["audio_s_per_pass_per_gpu"] * world}
        ms_step, extra = dt * 1e3, {}
    else:
        probe = cpu_sample_time(c, nfe_timed=1, threads=threads, warm=1)
        budget = 150.0
        n = int(max(1, min(c["nfe"], (budget / (K + min(W, 1)) - probe["t_text"] - probe["t_voc"]) / probe["t_nfe"])))
        steps = []
        for _ in range(min(W, 1) + K):
            t0 = time.perf_counter()
            r = cpu_sample_time(c, nfe_timed=n, threads=threads, warm=0)
            steps.append((time.perf_counter() - t0, r))
        steps = steps[min(W, 1):]
        t_nfe = statistics.mean(r["t_nfe"] for _, r in steps)
        t_fix = statistics.mean(r["t_text"] + r["t_voc"] for _, r in steps)
        total = t_fix + c["nfe"] * t_nfe
        value = probe["audio_s"] / total
        ms_step = total * 1e3
        sample = (f"each step = ONE utterance of the batch (B=1{' -> CFG batch 2' if c['kind'] != 'distill' else ''}) "
                  f"through text encoder + {n} of {c['nfe']} NFEs of the Euler loop + vocoder on the torch-fp32 CPU "
                  f"oracle port (kind: port -- the reference is Python and cannot travel to the GPU box); ms_per_step and "
                  f"value are scaled to the full {c['nfe']} NFEs")
        scale = {"utterances_timed_per_step": 1, "utterances_per_step_in_config": B * world, "nfe_timed_per_step": n,
                 "nfe_per_sample": c["nfe"]}
        extra = {"ms_per_nfe": t_nfe * 1e3}
    line = {
        "impl": "reference", "metric": "audio_sec_per_sec", "value": value, "unit": "audio-s/s",
        "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfgd,
        "sample_scale": scale,
        "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    line.update(extra)
    _emit(line)


# ------------------------------------------------------------------------------------------------ GPU arm
def residual_bytes(model_cfg: dict, nb: int, T: int, nfe: int) -> float:
    """Algorithmic HBM bytes of the residual-GEMM class for one sample() call: per output row the A operand (K bf16)
    + the residual stream read and written in its s24 storage format (bf16 hi plane + byte lo plane = 3 bytes per
    element each way; the hi plane is the next GEMM's A operand, there is no separate bf16 copy) = 2 K + 6 D, plus
    3 D for the layer input the second feed-forward's bypass reads -- DESIGN.md 3.1."""
    D, Fd = model_cfg["fm_decoder_dim"], model_cfg["fm_decoder_feedforward_dim"]
    ks = [Fd * 3 // 4, Fd, Fd * 5 // 4, 3 * D // 4, D + 64, D + 64]
    by = 0.0
    for ds, nl in zip(model_cfg["fm_decoder_downsampling_factor"], model_cfg["fm_decoder_num_layers"]):
        rows = nb * ((T + ds - 1) // ds)
        by += nl * sum(rows * (kk * 2.0 + D * (9.0 if i == 1 else 6.0)) for i, kk in enumerate(ks))
    return by * nfe


def attention_exps(model_cfg: dict, nb: int, T: int, nfe: int):
    """Softmax exponentials of one sample() call, the irreducible MUFU work of the attention kernels: per layer the two
    SelfAttention launches exponentiate H x L x L scores per sequence, the NonlinAttention launch head 0's L x L."""
    H = model_cfg["fm_decoder_num_heads"]
    pairs = 0.0
    for ds, nl in zip(model_cfg["fm_decoder_downsampling_factor"], model_cfg["fm_decoder_num_layers"]):
        L = (T + ds - 1) // ds
        pairs += nl * float(nb) * L * L
    return {"attn_self": 2.0 * H * pairs * nfe, "attn_nonlin": pairs * nfe}


MUFU_PER_CLK_PER_SM, NUM_SMS = 16.0, 148   # measured: tools/micro/mufu_f16_rate.cu (fp32, f16x2 and bf16x2 alike)

TRAFFIC_PREFIX = {"gemm_residual": ["gemm_tc_kernel<7,", "gemm_tc_kernel<4,"], "gemm_swooshl": ["gemm_tc_kernel<1,"],
                  "gemm_glu": ["gemm_tc_kernel<2,"], "gemm_nonlin": ["gemm_tc_kernel<3,"],
                  "gemm_gelu": ["gemm_tc_kernel<6,"], "attn_self": ["attn_sa_kernel"],
                  "attn_nonlin": ["attn_na_tc_kernel"], "dwconv_swooshr": ["dwconv_kernel"],
                  "fbank": ["fbank_kernel"], "vocoder_elem": ["vc_"]}


def ncu_traffic(kernel_class: str):
    """DRAM bytes per launch of the dominant kernel from the latest committed `ncu --set full` capture
    (profiles/r*_traffic.json, keys = demangled kernel names; matched by prefix so that added template
    arguments do not break the lookup)."""
    # newest set = highest pass tag (r1e < r1k < r2j < r2r ...): file times are meaningless after a checkout
    tfiles = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")), key=os.path.basename)
    prefixes = TRAFFIC_PREFIX.get(kernel_class, [])
    for tpath in reversed(tfiles):
        tj = json.load(open(tpath))
        sel = [v for kk, v in tj.items() if any(kk.startswith(p) for p in prefixes)]
        n = sum(x["captured_launches"] for x in sel)
        if n:
            tag = os.path.basename(tpath)[:-len("_traffic.json")]
            by = sum((x["dram_read_bytes_per_launch"] + x["dram_write_bytes_per_launch"]) * x["captured_launches"]
                     for x in sel) / n
            return round(by), (f"bytes per launch, dram__bytes_read.sum + dram__bytes_write.sum, mean of {n} launches of "
                               f"{'/'.join(prefixes)}* in one `ncu --set full` capture of the config-2 step "
                               f"(profiles/{tag}_traffic.json, profiles/{tag}_ncu_full_summary.md); the captured launches "
                               f"lie in a downsampled stack, so compare with the algorithmic bytes of THOSE launches in "
                               f"the summary, not with the class mean")
    return None, None


def kernel_table(table, pk, res_bytes=None, exps=None, sm_mhz=None):
    tot_ms = sum(r["ms"] for r in table) or 1.0
    kernels = []
    for r in sorted(table, key=lambda r: -r["ms"]):
        ach = r["work"] / (r["ms"] * 1e-3) / (1e12 if r["tensor"] else 1e9) if r["ms"] > 0 else 0.0
        peak = pk["tf_sustained"] if r["tensor"] else pk["hbm"]
        row = {"name": r["name"], "bound": "tensor" if r["tensor"] else "hbm", "launches": r["launches"],
               "ms": round(r["ms"], 3), "share": round(r["ms"] / tot_ms, 4), "achieved": round(ach, 1),
               "unit": "TFLOP/s" if r["tensor"] else "GB/s", "frac": round(ach / peak, 4)}
        if r["name"] == "gemm_residual" and res_bytes and r["ms"] > 0:
            # every residual GEMM of this model is HBM-bound at the measured peaks (fp32 stream through the epilogue):
            # the class is reported against the roofline that binds it, the tensor view is kept beside it
            hb = res_bytes / (r["ms"] * 1e-3) / 1e9
            row["tensor_view"] = {"achieved": row["achieved"], "peak": peak, "unit": "TFLOP/s", "frac": row["frac"]}
            row.update({"bound": "hbm", "achieved": round(hb, 1), "unit": "GB/s", "frac": round(hb / pk["hbm"], 4),
                        "algorithmic_bytes_per_launch": round(res_bytes / max(1, r["launches"]))})
        if exps and r["name"] in exps and r["ms"] > 0 and sm_mhz:
            # the attention kernels do little tensor work by construction (dq = 32, dv = 12 / 384): what bounds them from
            # below is the softmax exponential on the MUFU pipe, 16 results / clk / SM at the clock the step ran at
            ach = exps[r["name"]] / (r["ms"] * 1e-3) / 1e9
            peak = MUFU_PER_CLK_PER_SM * NUM_SMS * sm_mhz * 1e6 / 1e9
            row["mufu_view"] = {"achieved": round(ach, 1), "peak": round(peak, 1), "unit": "Gexp/s",
                                "frac": round(ach / peak, 4), "sm_mhz": sm_mhz}
        kernels.append(row)
    return kernels


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS))
    ap.add_argument("--batch", type=int, default=0, help="utterances per GPU (default: the config's)")
    ap.add_argument("--ragged", action="store_true", help="configs 2/3: targets of 6-10 s, length-balanced shards")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-trace", type=int, default=0, help="diagnostic: N extra e2e steps with per-stage syncs")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--gemm-impl", type=int, default=0, help="diagnostic: 0 auto, 2 single-CTA tiles, 3 CTA pairs")
    ap.add_argument("--profile-out", default=None, help="write the per-kernel-class table (JSON) here")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from zipvoice_b200 import parallel, synth
    from zipvoice_b200.model import ZipVoiceB200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs CUDA devices; there is no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    W, K = max(3, args.warmup), max(1, args.steps)
    k, c = args.config, CONFIGS[args.config]
    pk = peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_steps(fn):
        """W warm-ups, then K steps between CUDA events with a barrier + synchronize on both sides; max over ranks."""
        for _ in range(W):
            fn()
        barrier()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        eng.launch_count_reset()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(K):
            fn()
        e1.record()
        barrier()
        launches = eng.launch_count()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        clocks = sampler.stop() if rank == 0 else None
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / K, launches, clocks

    def timed_e2e(fn):
        for _ in range(max(1, min(W, 2))):
            fn()
        barrier()
        t0 = time.perf_counter()
        marks = []
        for _ in range(K):
            fn()
            marks.append(time.perf_counter())
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], device=dev)
        if rank == 0:
            print("e2e step wall ms:", [round((b - a) * 1e3, 1) for a, b in zip([t0] + marks[:-1], marks)],
                  file=sys.stderr)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        return float(dt.item()) / K

    e2e = None
    if c["kind"] == "audio":
        # ------------------------------------------------------------ config 5: mel front-end + vocoder sweep
        from zipvoice_b200.audio import audio_engine
        eng = audio_engine(synth.make_vocos_state_dict(0), device=dev)
        B = 0
        cases = []
        for secs, b in AUDIO_SWEEP:
            wav_host = synth.make_audio(b, secs * 24000, seed=7 + rank).pin_memory()
            cases.append((secs, b, wav_host, wav_host.to(dev)))
        audio_s = world * sum(s * b for s, b in AUDIO_SWEEP)

        def step_device():
            for _, _, _, wav in cases:
                mel = eng.fbank(wav)
                eng.vocos_decode(mel.permute(0, 2, 1))

        ms_step, launches, clocks = timed_steps(step_device)
        if not args.no_e2e:
            outs = [torch.empty(b, 256 * ((s * 24000 + 128) // 256 - 1)).pin_memory() for s, b, _, _ in cases]

            def step_e2e():
                for (secs, b, wav_host, _), out in zip(cases, outs):
                    mel = eng.fbank(wav_host.to(dev, non_blocking=True))
                    wav = eng.vocos_decode(mel.permute(0, 2, 1))
                    if world > 1:
                        lens = torch.full((b,), mel.shape[1], dtype=torch.int64, device=dev)
                        parallel.gather_ragged(mel, lens)
                        parallel.gather_ragged(wav, lens * 0 + wav.shape[1])
                    out.copy_(wav, non_blocking=True)
                torch.cuda.synchronize()

            s_e2e = timed_e2e(step_e2e)
            e2e = {"value": audio_s / s_e2e, "unit": "audio-s/s",
                   "h2d_bytes_per_step": int(sum(x[2].numel() * 4 for x in cases)),
                   "d2h_bytes_per_step": int(sum(o.numel() * 4 for o in outs)),
                   "api": "Engine.fbank (VocosFbank.extract) + Engine.vocos_decode (Vocos.decode, seeded random vocoder "
                          "weights, parity of the vocoder oracle unpinned) per sweep case; N>1: NCCL gather of mels and "
                          "waveforms to rank 0"}
        nfe, res_bytes = 1, None
    else:
        # ------------------------------------------------------------ configs 1-4: the sampling loop
        kind = c["kind"]
        B = args.batch or c["B"]
        cfg = synth.BASE_MODEL_CFG
        sd = synth.make_state_dict(cfg, seed=0, distill=(kind == "distill"), dialog=(kind == "dialog"))
        sd.update({"vocoder." + kk: v for kk, v in synth.make_vocos_state_dict(0).items()})  # same handle: model + vocoder
        model = ZipVoiceB200(cfg, sd, synth.VOCAB_SIZE, synth.PAD_ID, kind=kind, device=dev)
        eng = model.engine
        if args.gemm_impl:
            eng.set_option("gemm_impl", args.gemm_impl)
        nfe = c["nfe"]
        P = c["prompt_frames"]

        # the job's world * B utterances; ragged: dealt to the ranks by length (same list on every rank)
        all_targets = target_lengths(c, world * B, args.ragged)
        mine = parallel.shard_indices(all_targets, world, rank) if args.ragged else list(range(rank * B, rank * B + B))
        assert len(mine) == B
        targets = [all_targets[i] for i in mine]
        T = P + max(targets)
        tokens, prompt_tokens = synth.make_tokens(B, c["n_prompt_tok"], c["n_text_tok"], seed=1234 + rank,
                                                  dialog=(kind == "dialog"))
        pf_host, pfl = synth.make_prompt_features(B, P, seed=4321 + rank)
        x0_host = synth.make_noise(B, T, seed=666 + rank)
        pf_host, x0_host = pf_host.pin_memory(), x0_host.pin_memory()
        features_lens = torch.tensor(targets, dtype=torch.int64)

        # device-resident inputs for `value`: conditions prepared once through the library
        text_cond, pad_mask = model.forward_text_inference_gt_duration(tokens, features_lens.to(dev), prompt_tokens,
                                                                       pfl.to(dev))
        speech = torch.zeros(B, T, 100, device=dev)
        speech[:, :P] = pf_host.to(dev)
        x0 = x0_host.to(dev)
        torch.cuda.synchronize()
        audio_s = sum(all_targets) * FRAME_S

        def step_device():
            return eng.sample(x0, text_cond, speech, pad_mask, num_step=nfe, guidance_scale=c["guidance"],
                              t_shift=c["t_shift"])

        ms_step, launches, clocks = timed_steps(step_device)

        # ---------------- end to end through the public API from pinned host buffers
        if not args.no_e2e:
            n_wav = 256 * (max(targets) - 1)
            out_host = torch.empty(B, n_wav).pin_memory()
            wav_lens = (features_lens.to(dev) - 1) * 256

            def sample_e2e():
                return model.sample(tokens, prompt_tokens, pf_host.to(dev, non_blocking=True), pfl,
                                    features_lens=features_lens, duration="real", num_step=nfe,
                                    guidance_scale=c["guidance"], t_shift=c["t_shift"],
                                    x0=x0_host.to(dev, non_blocking=True))

            def step_e2e():
                x1_wo, wo_lens, _, _ = sample_e2e()
                # bin/infer_zipvoice.py:374-378: undo the feature scale, vocode, clamp
                wav = eng.vocos_decode(x1_wo.permute(0, 2, 1) / 0.1, wo_lens).clamp(-1, 1)
                if world > 1:
                    parallel.gather_ragged(x1_wo, wo_lens)
                    parallel.gather_ragged(wav, wav_lens)
                out_host.copy_(wav, non_blocking=True)
                torch.cuda.synchronize()

            s_e2e = timed_e2e(step_e2e)
            if args.e2e_trace and rank == 0:  # diagnostic only: stage times with a sync after each stage
                for _ in range(args.e2e_trace):
                    w0 = time.perf_counter()
                    x1_wo, wo_lens, _, _ = sample_e2e()
                    torch.cuda.synchronize(); w1 = time.perf_counter()
                    wav = eng.vocos_decode(x1_wo.permute(0, 2, 1) / 0.1, wo_lens).clamp(-1, 1)
                    torch.cuda.synchronize(); w2 = time.perf_counter()
                    out_host.copy_(wav, non_blocking=True)
                    torch.cuda.synchronize(); w3 = time.perf_counter()
                    print("e2e trace ms: sample %.1f vocoder %.1f d2h %.1f" % ((w1 - w0) * 1e3, (w2 - w1) * 1e3,
                                                                               (w3 - w2) * 1e3), file=sys.stderr)
            n_tok = sum(len(a) + len(b) + 1 for a, b in zip(tokens, prompt_tokens))
            e2e = {"value": audio_s / s_e2e, "unit": "audio-s/s",
                   "h2d_bytes_per_step": int(pf_host.numel() * 4 + x0_host.numel() * 4 + n_tok * 8 + 3 * B * 8),
                   "d2h_bytes_per_step": int(out_host.numel() * 4),
                   "api": f"ZipVoiceB200.sample (text encoder + condition + {nfe}-NFE solver + prompt strip) + Vocos "
                          "decode to 24 kHz waveforms (seeded random vocoder weights, parity of the vocoder oracle "
                          "unpinned); N>1: NCCL gather_ragged of the mels and the waveforms to rank 0"}
        nb = B if kind == "distill" else 2 * B
        res_bytes = residual_bytes(cfg, nb, T, nfe)

    value = audio_s / (ms_step * 1e-3)

    # ---------------- instrumented step: per-kernel-class device time (CUDA events inside the library)
    eng.prof_enable(True)
    step_device()
    table = eng.prof_read()
    eng.prof_enable(False)
    exps = attention_exps(cfg, nb, T, nfe) if c["kind"] != "audio" else None
    clk = (clocks or {}).get("sm_mhz") or (clocks or {}).get("sm_max_mhz")
    if world > 1:  # the clock sampler runs on rank 0 only
        ct = torch.tensor([float(clk or 0.0)], device=dev)
        dist.broadcast(ct, src=0)
        clk = float(ct.item()) or None
    kernels = kernel_table(table, pk, res_bytes, exps, clk)
    top = kernels[0]
    roofline = {"kernel": top["name"], "bound": top["bound"], "achieved": top["achieved"],
                "peak": pk["tf_sustained"] if top["bound"] == "tensor" else pk["hbm"], "unit": top["unit"],
                "frac": top["frac"], "traffic": None, "peak_source": pk["source"] + ", sustained figure",
                "avg_launch_ms": round(top["ms"] / max(1, top["launches"]), 4)}
    for kk in ("algorithmic_bytes_per_launch", "tensor_view"):
        if kk in top:
            roofline[kk] = top[kk]
    traffic, note = ncu_traffic(top["name"])
    if traffic is not None:
        roofline["traffic"], roofline["traffic_note"] = traffic, note
    flops_step = sum(r["work"] for r in table if r["tensor"])
    whole = {"algorithmic_tflop_per_nfe": round(flops_step / nfe / 1e12, 3),
             "tensor_frac_whole_step": round(flops_step / (ms_step * 1e-3) / 1e12 / pk["tf_sustained"], 4)}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_obj(k, c, os.cpu_count() or 1)
    line = {
        "metric": "audio_sec_per_sec", "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_step, "ms_per_nfe": ms_step / nfe, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16" if c["kind"] != "audio" else "f32 (FFT / iSTFT) + bf16 (vocoder GEMMs)",
        "data": "synthetic", "config": config_dict(k, c, B, world, args.ragged),
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "whole_step": whole,
        "cpu_baseline": cpu, "kernels": kernels,
    }
    if c["kind"] == "audio":
        line.pop("ms_per_nfe")
    if args.profile_out:
        os.makedirs(os.path.dirname(os.path.abspath(args.profile_out)), exist_ok=True)
        json.dump({"ms_per_step": ms_step, "kernels": kernels, "whole_step": whole}, open(args.profile_out, "w"),
                  indent=1)
    _emit(line)
    if world > 1:
        dist.destroy_process_group()


class _StdoutToStderr:
    """Everything but the result line goes to stderr: libraries write to file descriptor 1 behind Python's back
    (NCCL prints its version banner there), and the driver expects exactly one JSON line on stdout."""

    def __enter__(self):
        sys.stdout.flush()
        self._saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def restore(self):
        sys.stdout.flush()
        os.dup2(self._saved, 1)

    def __exit__(self, *exc):
        self.restore()
        os.close(self._saved)
        return False


_REDIRECT = None


def _emit(line: dict) -> None:
    """Print the result line on the real stdout."""
    if _REDIRECT is not None:
        _REDIRECT.restore()
    print(json.dumps(line), flush=True)
    if _REDIRECT is not None:
        os.dup2(2, 1)


if __name__ == "__main__":
    with _StdoutToStderr() as _REDIRECT:
        main()

#!/usr/bin/env python
"""Benchmark of the ZipVoice sampling hot path on B200 (contract: see DESIGN.md "Measurement").

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # CPU reference arm (oracle port, host cores)

One "step" = one full pass of the hot path over one batch: ``EulerSolver.sample`` with 16 NFE and
classifier-free guidance on B = 64 utterances per GPU (3 s prompt + 10 s target, T = 1219 frames of
100-bin 24 kHz mel) of the 123 M ZipVoice fm_decoder -- BASELINE.json configs[1].  Metric:
generated audio-seconds per second (prompt frames excluded), whole job over all ranks; also ms per NFE.

  value : inputs resident in HBM, timed with CUDA events around K steps (barrier + sync both sides,
          max over ranks).
  e2e   : the same batch through the public API ``ZipVoiceB200.sample`` (text encoder, text condition,
          prompt padding, solver, prompt strip) + the vocoder, from pinned HOST buffers to HOST waveforms,
          H2D and D2H copies inside the timed region; N > 1 adds the NCCL gather of the waveforms to rank 0.
  roofline / kernels : one extra instrumented step in which the library brackets every launch of its
          kernels with CUDA events on the launching stream (zv_prof_*).
  cpu_baseline : the CPU oracle (torch fp32 restatement of the reference, oracle/) timed on the host
          cores on a bounded sample of the same workload (rank 0, N = 1 only).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FRAME_S = 256.0 / 24000.0
PROMPT_FRAMES = 281      # 3 s
TARGET_FRAMES = 938      # 10 s
NUM_STEP, GUIDANCE, T_SHIFT = 16, 1.0, 0.5
WORKLOAD = ("configs[1]: ZipVoice 123M random-init, B=64 per GPU, 3 s prompt + 10 s target (T=1219, 100-bin "
            "24 kHz mel), 16 NFE Euler, CFG guidance 1.0 (batch doubled to 128), t_shift 0.5")


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


def cpu_reference_sample(nfe_reps: int, threads: int):
    """Times the CPU oracle (torch fp32 port of the reference) on a bounded sample of the workload:
    ONE utterance of the batch (B=1 -> CFG batch 2, T=1219), `nfe_reps` guided velocity evaluations;
    a 16-NFE sample of that utterance is 16x the median.  Returns (audio_s_per_s, seconds_per_nfe)."""
    import torch
    from oracle import zipvoice_oracle as O
    from tests.cases import case_inputs
    from zipvoice_b200 import synth
    torch.set_num_threads(threads)
    cfg = synth.BASE_MODEL_CFG
    sd = synth.make_state_dict(cfg, seed=0)
    T = PROMPT_FRAMES + TARGET_FRAMES
    x0, text, speech, mask = case_inputs(1, T, PROMPT_FRAMES, seed=21, ragged=False)
    ts = []
    with torch.inference_mode():
        for i in range(nfe_reps + 1):
            t0 = time.perf_counter()
            O.cfg_velocity(sd, cfg, torch.tensor(0.3), x0, text, speech, mask, GUIDANCE)
            ts.append(time.perf_counter() - t0)
    per_nfe = statistics.median(ts[1:]) if nfe_reps > 0 else ts[0]
    return TARGET_FRAMES * FRAME_S / (NUM_STEP * per_nfe), per_nfe


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    K = max(1, args.steps)
    # each "step" = one guided NFE of one utterance (B=1, T=1219); warmup steps likewise
    import torch
    from oracle import zipvoice_oracle as O
    from tests.cases import case_inputs
    from zipvoice_b200 import synth
    torch.set_num_threads(threads)
    cfg = synth.BASE_MODEL_CFG
    sd = synth.make_state_dict(cfg, seed=0)
    T = PROMPT_FRAMES + TARGET_FRAMES
    x0, text, speech, mask = case_inputs(1, T, PROMPT_FRAMES, seed=21, ragged=False)
    with torch.inference_mode():
        for _ in range(max(1, min(args.warmup, 2))):
            O.cfg_velocity(sd, cfg, torch.tensor(0.3), x0, text, speech, mask, GUIDANCE)
        t0 = time.perf_counter()
        for _ in range(K):
            O.cfg_velocity(sd, cfg, torch.tensor(0.3), x0, text, speech, mask, GUIDANCE)
        per_nfe = (time.perf_counter() - t0) / K
    value = TARGET_FRAMES * FRAME_S / (NUM_STEP * per_nfe)
    sample = ("1 utterance of the batch (B=1 -> CFG batch 2, T=1219), each step = 1 guided fm_decoder NFE of the "
              "torch-fp32 CPU oracle port; audio-s/s = 10 s / (16 NFE x s/NFE)")
    _emit({
        "impl": "reference", "metric": "audio_sec_per_sec", "value": value, "unit": "audio-s/s",
        "n_gpus": args.gpus, "steps": K, "warmup": args.warmup, "ms_per_step": per_nfe * 1e3,
        "ms_per_nfe": per_nfe * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": {"workload": WORKLOAD},
        "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="utterances per GPU")
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
    from zipvoice_b200 import synth
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

    cfg = synth.BASE_MODEL_CFG
    sd = synth.make_state_dict(cfg, seed=0)
    sd.update({"vocoder." + k: v for k, v in synth.make_vocos_state_dict(0).items()})  # same handle: model + vocoder
    model = ZipVoiceB200(cfg, sd, synth.VOCAB_SIZE, synth.PAD_ID, kind="zipvoice", device=dev)
    eng = model.engine
    if args.gemm_impl:
        eng.set_option("gemm_impl", args.gemm_impl)
    B, T = args.batch, PROMPT_FRAMES + TARGET_FRAMES

    # ---------------- synthetic utterances (different per rank: data-parallel shards)
    tokens, prompt_tokens = synth.make_tokens(B, 45, 150, seed=1234 + rank)
    pf_host, pfl = synth.make_prompt_features(B, PROMPT_FRAMES, seed=4321 + rank)
    x0_host = synth.make_noise(B, T, seed=666 + rank)
    pf_host, x0_host = pf_host.pin_memory(), x0_host.pin_memory()
    features_lens = torch.full((B,), TARGET_FRAMES, dtype=torch.int64)

    # device-resident inputs for `value`: conditions prepared once through the library
    text_cond, pad_mask = model.forward_text_inference_gt_duration(tokens, features_lens.to(dev), prompt_tokens,
                                                                   pfl.to(dev))
    speech = torch.zeros(B, T, 100, device=dev)
    speech[:, :PROMPT_FRAMES] = pf_host.to(dev)
    x0 = x0_host.to(dev)
    torch.cuda.synchronize()

    def step_device():
        return eng.sample(x0, text_cond, speech, pad_mask, num_step=NUM_STEP, guidance_scale=GUIDANCE,
                          t_shift=T_SHIFT)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(W):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    eng.launch_count_reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step_device()
    e1.record()
    barrier()
    launches = eng.launch_count()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms_total], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / K
    audio_s = world * B * TARGET_FRAMES * FRAME_S
    value = audio_s / (ms_step * 1e-3)

    # ---------------- end to end through the public API from pinned host buffers
    e2e = None
    if not args.no_e2e:
        n_wav = 256 * (TARGET_FRAMES - 1)
        out_host = torch.empty(B, n_wav).pin_memory()
        gathered = [torch.empty(B, n_wav, device=dev) for _ in range(world)] if (world > 1 and rank == 0) else None

        def step_e2e():
            x1_wo, wo_lens, _, _ = model.sample(tokens, prompt_tokens, pf_host.to(dev, non_blocking=True), pfl,
                                                features_lens=features_lens, duration="real", num_step=NUM_STEP,
                                                guidance_scale=GUIDANCE, t_shift=T_SHIFT,
                                                x0=x0_host.to(dev, non_blocking=True))
            # bin/infer_zipvoice.py:374-378: undo the feature scale, vocode, clamp
            wav = eng.vocos_decode(x1_wo.permute(0, 2, 1) / 0.1).clamp(-1, 1)
            if world > 1:
                dist.gather(wav, gathered, dst=0)
            out_host.copy_(wav, non_blocking=True)
            torch.cuda.synchronize()

        for _ in range(max(1, min(W, 2))):
            step_e2e()
        barrier()
        t0 = time.perf_counter()
        marks = []
        for _ in range(K):
            step_e2e()
            marks.append(time.perf_counter())
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], device=dev)
        if rank == 0:
            print("e2e step wall ms:", [round((b - a) * 1e3, 1) for a, b in zip([t0] + marks[:-1], marks)],
                  file=sys.stderr)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        if args.e2e_trace and rank == 0:  # diagnostic only: stage times with a sync after each stage
            for _ in range(args.e2e_trace):
                w0 = time.perf_counter()
                x1_wo, _, _, _ = model.sample(tokens, prompt_tokens, pf_host.to(dev, non_blocking=True), pfl,
                                              features_lens=features_lens, duration="real", num_step=NUM_STEP,
                                              guidance_scale=GUIDANCE, t_shift=T_SHIFT,
                                              x0=x0_host.to(dev, non_blocking=True))
                torch.cuda.synchronize(); w1 = time.perf_counter()
                wav = eng.vocos_decode(x1_wo.permute(0, 2, 1) / 0.1).clamp(-1, 1)
                torch.cuda.synchronize(); w2 = time.perf_counter()
                out_host.copy_(wav, non_blocking=True)
                torch.cuda.synchronize(); w3 = time.perf_counter()
                print("e2e trace ms: sample %.1f vocoder %.1f d2h %.1f" % ((w1 - w0) * 1e3, (w2 - w1) * 1e3,
                                                                           (w3 - w2) * 1e3), file=sys.stderr)
        n_tok = sum(len(a) + len(b) + 1 for a, b in zip(tokens, prompt_tokens))
        e2e = {"value": audio_s / (float(dt.item()) / K), "unit": "audio-s/s",
               "h2d_bytes_per_step": int(pf_host.numel() * 4 + x0_host.numel() * 4 + n_tok * 8 + 3 * B * 8),
               "d2h_bytes_per_step": int(out_host.numel() * 4),
               "api": "ZipVoiceB200.sample (text encoder + condition + 16-NFE CFG solver + prompt strip) + "
                      "Vocos decode to 24 kHz waveforms (seeded random vocoder weights); N>1: NCCL gather of the "
                      "waveforms to rank 0"}

    # ---------------- instrumented step: per-kernel-class device time (CUDA events inside the library)
    eng.prof_enable(True)
    step_device()
    table = eng.prof_read()
    eng.prof_enable(False)
    pk = peaks()
    tot_ms = sum(r["ms"] for r in table) or 1.0
    kernels = []
    for r in sorted(table, key=lambda r: -r["ms"]):
        ach = r["work"] / (r["ms"] * 1e-3) / (1e12 if r["tensor"] else 1e9) if r["ms"] > 0 else 0.0
        peak = pk["tf_sustained"] if r["tensor"] else pk["hbm"]
        kernels.append({"name": r["name"], "bound": "tensor" if r["tensor"] else "hbm", "launches": r["launches"],
                        "ms": round(r["ms"], 3), "share": round(r["ms"] / tot_ms, 4), "achieved": round(ach, 1),
                        "unit": "TFLOP/s" if r["tensor"] else "GB/s", "frac": round(ach / peak, 4)})
    top = kernels[0]
    roofline = {"kernel": top["name"], "bound": top["bound"], "achieved": top["achieved"],
                "peak": pk["tf_sustained"] if top["bound"] == "tensor" else pk["hbm"], "unit": top["unit"],
                "frac": top["frac"], "traffic": None, "peak_source": pk["source"] + ", sustained figure",
                "avg_launch_ms": round(top["ms"] / max(1, top["launches"]), 4)}
    # ncu DRAM traffic of the dominant kernel (one `ncu --set full` capture, committed under profiles/)
    import glob
    tfiles = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")))
    tpath = tfiles[-1] if tfiles else ""
    ttag = os.path.basename(tpath)[:-len("_traffic.json")] if tpath else ""
    knames = {"gemm_residual": ["gemm_tc_kernel<4, 1>", "gemm_tc_kernel<4, 0>"], "gemm_swooshl": ["gemm_tc_kernel<1, 0>"],
              "attn_self": ["attn_sa_kernel"], "attn_nonlin": ["attn_na_tc_kernel"]}.get(top["name"], [])
    if tpath and knames:
        tj = json.load(open(tpath))
        sel = [tj[k] for k in knames if k in tj]
        n = sum(x["captured_launches"] for x in sel)
        if n:
            roofline["traffic"] = round(sum((x["dram_read_bytes_per_launch"] + x["dram_write_bytes_per_launch"]) *
                                            x["captured_launches"] for x in sel) / n)
            roofline["traffic_note"] = (f"bytes per launch, dram__bytes_read.sum + dram__bytes_write.sum, mean of {n} launches "
                                        f"captured in a downsampled stack (profiles/{ttag}_ncu_full_summary.md, "
                                        f"profiles/{ttag}_traffic.json); below the algorithmic bytes of those launches because "
                                        "part of the residual stream is still L2-resident from the producer kernel")
    if top["name"] == "gemm_residual":
        # The residual GEMMs (N = 512, K = 384 ... 1920) carry the fp32 residual stream through their epilogue and
        # are HBM-bound for every shape of this model (ncu: DRAM 65-76 % busy, tensor pipe 25-50 %; at the measured
        # peaks the byte time exceeds the flop time for all K <= 1920), so the binding roofline is HBM.  Algorithmic
        # bytes per output row = A operand (K bf16) + fp32 stream read + fp32 stream write + bf16 shadow write
        # = 2 K + 10 D  (2 K + 8 D for the last feed-forward, whose shadow nobody reads).  The tensor-pipe view of the same time is kept beside it.
        cfgm = synth.BASE_MODEL_CFG
        D, Fd = cfgm["fm_decoder_dim"], cfgm["fm_decoder_feedforward_dim"]
        ks = [Fd * 3 // 4, Fd, Fd * 5 // 4, 3 * D // 4, D + 64, D + 64]
        by = 0.0
        for ds, nl in zip(cfgm["fm_decoder_downsampling_factor"], cfgm["fm_decoder_num_layers"]):
            rows = 2 * B * ((T + ds - 1) // ds)
            # the last feed-forward (ks[2]) writes no bf16 shadow: BiasNorm reads the fp32 stream
            by += nl * sum(rows * (k * 2.0 + D * (8.0 if i == 2 else 10.0)) for i, k in enumerate(ks))
        by *= NUM_STEP
        tensor_view = {"achieved": roofline["achieved"], "peak": roofline["peak"], "unit": roofline["unit"],
                       "frac": roofline["frac"]}
        ach = by / (top["ms"] * 1e-3) / 1e9
        roofline.update({"bound": "hbm", "achieved": round(ach, 1), "peak": pk["hbm"], "unit": "GB/s",
                         "frac": round(ach / pk["hbm"], 4), "algorithmic_bytes_per_launch": round(by / top["launches"]),
                         "tensor_view": tensor_view})
    flops_step = sum(r["work"] for r in table if r["tensor"])
    whole = {"algorithmic_tflop_per_nfe": round(flops_step / NUM_STEP / 1e12, 3),
             "tensor_frac_whole_step": round(flops_step / (ms_step * 1e-3) / 1e12 / pk["tf_sustained"], 4)}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, per_nfe = cpu_reference_sample(nfe_reps=3, threads=threads)
        cpu = {"value": v, "unit": "audio-s/s", "cores": threads, "kind": "port", "s_per_nfe": per_nfe,
               "sample": "1 utterance (B=1 -> CFG batch 2, T=1219), 1 warm-up + 3 guided NFE of the torch-fp32 CPU "
                         "oracle port, median; audio-s/s = 10 s / (16 x s/NFE)"}
    line = {
        "metric": "audio_sec_per_sec", "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_step, "ms_per_nfe": ms_step / NUM_STEP, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch_per_gpu": B, "frames": T, "nfe": NUM_STEP, "guidance_scale": GUIDANCE,
                   "t_shift": T_SHIFT, "weights": "synth.make_state_dict(seed=0), 123M",
                   "l2": "activations of one NFE (several GB) exceed the 126 MB L2; no explicit flush needed",
                   "parallelism": f"dp{world} over utterances"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "whole_step": whole,
        "cpu_baseline": cpu, "kernels": kernels,
    }
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

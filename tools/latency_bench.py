#!/usr/bin/env python
"""Single-utterance latency (BASELINE.json configs[0]: B=1, 3 s prompt + 5 s target, 16 NFE, CFG) and small batches,
plain launches vs CUDA-graph replay of the sample loop.  usage: python tools/latency_bench.py [--out file.json]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from zipvoice_b200 import synth  # noqa: E402
from zipvoice_b200.model import ZipVoiceB200  # noqa: E402

dev = torch.device("cuda:0")
FRAME_S = 256.0 / 24000.0


def timed(fn, reps=5, warm=3):
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "latency.json"))
    args = ap.parse_args()
    rows = []
    for kind, nfe, guid in (("zipvoice", 16, 1.0), ("distill", 4, 3.0)):
        cfg = synth.BASE_MODEL_CFG
        sd = synth.make_state_dict(cfg, seed=0, distill=(kind == "distill"))
        model = ZipVoiceB200(cfg, sd, synth.VOCAB_SIZE, synth.PAD_ID, kind=kind, device=dev)
        eng = model.engine
        for B, prompt, target in ((1, 281, 469), (4, 281, 469), (8, 281, 938)):
            T = prompt + target
            g = torch.Generator().manual_seed(B)
            x0 = torch.randn(B, T, 100, generator=g).to(dev)
            text = (torch.randn(B, T, 100, generator=g) * 0.5).to(dev)
            speech = torch.zeros(B, T, 100)
            speech[:, :prompt] = torch.randn(B, prompt, 100, generator=g) * 0.5 - 0.5
            speech = speech.to(dev)
            mask = torch.zeros(B, T, dtype=torch.bool, device=dev)
            res = {"kind": kind, "B": B, "T": T, "nfe": nfe}
            outs = {}
            for name, rows_opt in (("plain", 0), ("graph", 1 << 20)):
                eng.set_option("graph_max_rows", rows_opt)
                fn = lambda: eng.sample(x0, text, speech, mask, num_step=nfe, guidance_scale=guid, t_shift=0.5)
                ms = timed(fn)
                outs[name] = fn().clone()
                res[name + "_ms_per_sample"] = ms
                res[name + "_ms_per_nfe"] = ms / nfe
                res[name + "_audio_s_per_s"] = B * target * FRAME_S / (ms * 1e-3)
            res["bit_identical"] = bool(torch.equal(outs["plain"], outs["graph"]))
            rows.append(res)
            print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in res.items()})
        eng.close()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()

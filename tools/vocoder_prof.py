#!/usr/bin/env python
"""Per-kernel-class device time of the mel front-end and the vocoder head (CUDA events inside the library).
usage: python tools/vocoder_prof.py [B] [seconds]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from zipvoice_b200 import synth  # noqa: E402
from zipvoice_b200.audio import VocosB200, VocosFbankB200  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 10.0
voc = VocosB200(synth.make_vocos_state_dict(0))
fb = VocosFbankB200(engine=voc.engine)
wav = synth.make_audio(B, int(secs * 24000), seed=7).cuda()
mel = fb.extract_batch(wav).permute(0, 2, 1).contiguous()  # (B, 100, T)
for _ in range(2):
    voc.decode(mel)
    fb.extract_batch(wav)
torch.cuda.synchronize()
eng = voc.engine
eng.prof_enable(True)
fb.extract_batch(wav)
voc.decode(mel)
torch.cuda.synchronize()
table = eng.prof_read()
eng.prof_enable(False)
tot = sum(r["ms"] for r in table)
print(f"B={B} clip={secs}s frames={mel.shape[2]} total {tot:.3f} ms")
for r in sorted(table, key=lambda r: -r["ms"]):
    if r["ms"] <= 0:
        continue
    ach = r["work"] / (r["ms"] * 1e-3) / (1e12 if r["tensor"] else 1e9)
    print(f"{r['name']:18s} launches {r['launches']:3d} {r['ms']:8.3f} ms {100 * r['ms'] / tot:5.1f} %  {ach:8.1f} "
          f"{'TFLOP/s' if r['tensor'] else 'GB/s'}")

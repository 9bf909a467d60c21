"""Per-class device time of one B=1 sample call (config 1)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import synth
from zipvoice_b200.model import ZipVoiceB200
dev = torch.device("cuda:0")
cfg = synth.BASE_MODEL_CFG
model = ZipVoiceB200(cfg, synth.make_state_dict(cfg, seed=0), synth.VOCAB_SIZE, synth.PAD_ID, device=dev)
eng = model.engine
B, T = int(sys.argv[1]) if len(sys.argv) > 1 else 1, 750
x0 = torch.randn(B, T, 100, device=dev); text = torch.randn(B, T, 100, device=dev) * 0.5
speech = torch.zeros(B, T, 100, device=dev); mask = torch.zeros(B, T, dtype=torch.bool, device=dev)
fn = lambda: eng.sample(x0, text, speech, mask, num_step=16, guidance_scale=1.0, t_shift=0.5)
for _ in range(3): fn()
eng.prof_enable(True); fn(); tab = eng.prof_read(); eng.prof_enable(False)
tot = sum(r["ms"] for r in tab)
print("total ms", round(tot, 2), "per NFE", round(tot / 16, 3))
for r in sorted(tab, key=lambda r: -r["ms"]):
    print(f'{r["name"]:18s} launches {r["launches"]:5d}  ms {r["ms"]:8.3f}  us/launch {1e3 * r["ms"] / max(1, r["launches"]):7.2f}')

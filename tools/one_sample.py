"""One B=1 sample call, plain launches then graph replay (debug aid).  usage: python tools/one_sample.py [B] [T]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zipvoice_b200 import synth
from zipvoice_b200.model import ZipVoiceB200
dev = torch.device("cuda:0")
cfg = synth.BASE_MODEL_CFG
model = ZipVoiceB200(cfg, synth.make_state_dict(cfg, seed=0), synth.VOCAB_SIZE, synth.PAD_ID, device=dev)
eng = model.engine
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
T = int(sys.argv[2]) if len(sys.argv) > 2 else 750
x0 = torch.randn(B, T, 100, device=dev); text = torch.randn(B, T, 100, device=dev) * 0.5
speech = torch.zeros(B, T, 100, device=dev); mask = torch.zeros(B, T, dtype=torch.bool, device=dev)
for mode in (0, 1):
    eng.set_option("graph_max_rows", (1 << 20) if mode else 0)
    out = eng.sample(x0, text, speech, mask, num_step=4, guidance_scale=1.0, t_shift=0.5)
    torch.cuda.synchronize()
    print("mode", mode, "ok", float(out.abs().mean()), flush=True)

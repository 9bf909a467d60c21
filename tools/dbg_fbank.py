import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from zipvoice_b200 import synth
from zipvoice_b200.audio import VocosFbankB200
from oracle import fbank_oracle as FO
fb = VocosFbankB200()
g = np.load("tests/golden/fbank.npz")
for name, n in (("a", 72000), ("b", 24077), ("c", 30200)):
    wav = synth.make_audio(1, n, seed=7 + n)
    m1 = fb.extract(wav, 24000).cpu(); m2 = fb.extract(wav, 24000).cpu()
    want = torch.from_numpy(g["mel_" + name]); orc = FO.vocos_fbank(wav[0])
    d = (m1 - want).abs()
    i = int(d.argmax()); t, b = divmod(i, 100)
    print(name, "rerun-equal", bool((m1 == m2).all()), "max", float(d.max()), "at frame", t, "bin", b, "val", float(want[t, b]),
          "oracle-vs-golden", float((orc - want).abs().max()), "mean", float(d.mean()))
    # 64-bit reference
    x = wav[0].double(); xp = torch.nn.functional.pad(x[None, None], (512, 512), mode="reflect")[0, 0]
    fr = xp.unfold(0, 1024, 256) * torch.hann_window(1024, periodic=True, dtype=torch.float64)
    mel64 = (torch.fft.rfft(fr).abs() @ torch.from_numpy(FO.mel_filterbank()).double()).clamp(min=1e-7).log()[: want.shape[0]]
    print("   vs fp64: cuda", float((m1.double() - mel64).abs().max()), "golden", float((want.double() - mel64).abs().max()))

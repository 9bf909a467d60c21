"""Parity of the mel front-end and the vocoder head (CUDA, through the C ABI) against the CPU oracles.
fbank: fp32 arithmetic end to end, compared with the reference golden (tests/golden/fbank.npz) and the
oracle; tolerance 2e-4 absolute on log-mel (fp32 FFT vs torch's).  Vocoder: bf16 tensor-core GEMMs vs
the fp32 oracle restatement (parity UNPINNED against the third-party vocos, see oracle/vocos_oracle.py);
tolerance rel-L2 <= 2e-2 on the waveform, 1e-2 on the complex spectrum path checked through the wav."""
import os

import numpy as np
import pytest
import torch

from tests.cases import rel_l2
from tests.conftest import GOLDEN

pytestmark = pytest.mark.gpu

_cache = {}


def audio():
    from oracle import vocos_oracle as VO
    from zipvoice_b200.audio import VocosB200, VocosFbankB200
    if "a" not in _cache:
        sd = VO.make_vocos_state_dict(0)
        voc = VocosB200(sd)
        _cache["a"] = (VocosFbankB200(engine=voc.engine), voc, sd)
    return _cache["a"]


def test_fbank_reference_golden():
    from zipvoice_b200 import synth
    fb, _, _ = audio()
    g = np.load(os.path.join(GOLDEN, "fbank.npz"))
    for name in ("a", "b", "c"):
        wav = torch.from_numpy(g["wav_" + name])
        n = wav.shape[1]
        mel = fb.extract(wav, 24000).cpu()
        want = torch.from_numpy(g["mel_" + name])
        assert mel.shape == want.shape
        err = float((mel - want).abs().max())
        assert err < 2e-4, f"{name}: max-abs {err:.3e}"


def test_fbank_batch_ragged_and_numpy():
    from oracle import fbank_oracle as FO
    from zipvoice_b200 import synth
    fb, _, _ = audio()
    lens = torch.tensor([48000, 31234, 20000, 47999])
    wav = synth.make_audio(4, 48000, seed=3)
    wav = wav * (torch.arange(48000)[None] < lens[:, None])
    mel = fb.extract_batch(wav, lens).cpu()
    assert mel.shape == (4, FO.num_frames(48000), 100)
    for b in range(4):
        want = FO.vocos_fbank(wav[b, : int(lens[b])])
        t = want.shape[0]
        assert float((mel[b, :t] - want).abs().max()) < 2e-4
        assert float(mel[b, t:].abs().max()) == 0.0 if t < mel.shape[1] else True
    out = fb.extract(wav[0].numpy(), 24000)
    assert isinstance(out, np.ndarray) and out.shape == (188, 100)
    stereo = torch.stack([wav[0], wav[1]])
    m1 = fb.extract(stereo, 24000).cpu()
    want = FO.vocos_fbank(stereo.mean(0))
    assert float((m1 - want).abs().max()) < 2e-4


def test_bigvgan_fbank_reference_golden_and_ragged_batch():
    """BigVGANFbank features (feature.py:133-204) on the mel kernel: the golden of the unmodified reference class
    (tests/golden/fbank_bigvgan.npz; mono, stereo averaged, two-channel), then a ragged padded batch against the
    oracle.  log floor 1e-5 and the 1e-9 under the square root keep every value O(1): 2e-4 absolute as for VocosFbank."""
    from oracle import fbank_oracle as FO
    from zipvoice_b200 import synth
    from zipvoice_b200.audio import BigVGANFbankB200
    eng = audio()[0].engine
    fb = BigVGANFbankB200(engine=eng)
    g = np.load(os.path.join(GOLDEN, "fbank_bigvgan.npz"))
    for name in ("a", "b", "c", "d"):
        wav = torch.from_numpy(g["wav_" + name])
        mel = fb.extract(wav, 24000).cpu()
        want = torch.from_numpy(g["mel_" + name])
        assert mel.shape == want.shape
        err = float((mel - want).abs().max())
        assert err < 2e-4, f"{name}: max-abs {err:.3e}"
    st = torch.from_numpy(g["wav_stereo"])
    assert float((fb.extract(st, 24000).cpu() - torch.from_numpy(g["mel_stereo_mono"])).abs().max()) < 2e-4
    fb2 = BigVGANFbankB200(engine=eng, num_channels=2)
    m2 = fb2.extract(st, 24000).cpu()
    assert m2.shape == (FO.num_frames(st.shape[1]), 200)
    assert float((m2 - torch.from_numpy(g["mel_stereo_2ch"])).abs().max()) < 2e-4
    lens = torch.tensor([48000, 31234, 20000, 47999, 1024 + 128])
    wav = synth.make_audio(5, 48000, seed=13)
    wav = wav * (torch.arange(48000)[None] < lens[:, None])
    mel = fb.extract_batch(wav, lens).cpu()
    for b in range(5):
        want = FO.bigvgan_fbank(wav[b, : int(lens[b])])
        t = want.shape[0]
        assert float((mel[b, :t] - want).abs().max()) < 2e-4, b
        if t < mel.shape[1]:
            assert float(mel[b, t:].abs().max()) == 0.0
    with pytest.raises(AssertionError):
        fb.extract(wav[0], 16000)


@pytest.mark.parametrize("B,T", [(1, 50), (3, 281), (2, 938)])
def test_vocos_decode_vs_oracle(B, T):
    from oracle import vocos_oracle as VO
    _, voc, sd = audio()
    g = torch.Generator().manual_seed(100 + T)
    mel = torch.randn(B, 100, T, generator=g) * 0.8 - 0.5
    wav = voc.decode(mel).cpu()
    want = VO.decode(sd, mel)
    assert wav.shape == want.shape == (B, 256 * (T - 1))
    r = rel_l2(wav, want)
    assert r < 2e-2, f"waveform rel-L2 {r:.3e}"


def test_vocos_decode_ragged_equals_single():
    """Rows of a padded batch decode as if each utterance were decoded alone (lens masks the convs)."""
    from oracle import vocos_oracle as VO
    _, voc, sd = audio()
    g = torch.Generator().manual_seed(5)
    lens = torch.tensor([120, 77, 33])
    mel = torch.randn(3, 100, 120, generator=g) * 0.8 - 0.5
    wav = voc.decode(mel, lens).cpu()
    for b in range(3):
        n = int(lens[b])
        want = VO.decode(sd, mel[b : b + 1, :, :n])[0]
        r = rel_l2(wav[b, : 256 * (n - 1)], want)
        assert r < 2e-2, f"row {b}: rel-L2 {r:.3e}"
        assert float(wav[b, 256 * (n - 1):].abs().max() if n < 120 else 0.0) == 0.0


def test_fbank_vocoder_roundtrip_shapes():
    """mel -> wav -> mel plumbing at BASELINE config-5 like sizes (size-independent property: frame counts)."""
    fb, voc, _ = audio()
    mel = torch.randn(8, 100, 400) * 0.5 - 1.0
    wav = voc.decode(mel)
    assert wav.shape == (8, 256 * 399) and bool(torch.isfinite(wav).all())
    mel2 = fb.extract_batch(wav)
    assert mel2.shape == (8, 399, 100) and bool(torch.isfinite(mel2).all())


@pytest.mark.parametrize("sr", [16000, 22050, 44100, 48000])
def test_resample_torchaudio_golden(sr):
    """zv_resample against torchaudio.transforms.Resample(sr, 24000) outputs (tests/golden/resample.npz):
    same fp32 filter bank (computed in double), fp32 FMA dot products; 2e-5 absolute on |x| <= 1 signals."""
    from zipvoice_b200.audio import ResampleB200
    fb, voc, _ = audio()
    g = np.load(os.path.join(GOLDEN, "resample.npz"))
    x, want = torch.from_numpy(g[f"x_{sr}"]), torch.from_numpy(g[f"y_{sr}"])
    got = ResampleB200(sr, 24000, engine=fb.engine)(x).cpu()
    assert got.shape == want.shape
    assert float((got - want).abs().max()) < 2e-5


def test_resample_ragged_rows_equal_single_rows():
    """Rows of a padded batch with per-row lengths are resampled as if each had been handed over alone."""
    from oracle import resample_oracle as R
    from zipvoice_b200.audio import ResampleB200
    fb, voc, _ = audio()
    g = torch.Generator().manual_seed(5)
    lens = torch.tensor([7000, 4411, 1, 6999])
    x = torch.randn(4, 7000, generator=g) * 0.3
    rs = ResampleB200(44100, 24000, engine=fb.engine)
    got = rs(x, lens).cpu()
    for b in range(4):
        n = int(lens[b])
        want = torch.from_numpy(R.resample(x[b, :n].numpy(), 44100, 24000))
        assert float((got[b, : want.numel()] - want).abs().max()) < 2e-5
        assert float(got[b, want.numel():].abs().max() if want.numel() < got.shape[1] else 0.0) == 0.0
    same = ResampleB200(24000, 24000, engine=fb.engine)(x)
    assert same is x


def test_compute_fbank_bulk_equals_single_extraction(tmp_path):
    """Bulk extraction (bin/compute_fbank.py on the mel kernel): bucketed padded batches return exactly the
    per-recording features, in input order, and the npz holds the same arrays."""
    from zipvoice_b200.audio import compute_fbank
    fb, voc, _ = audio()
    g = torch.Generator().manual_seed(6)
    wavs = [torch.randn(n, generator=g) * 0.2 for n in (24000, 5000, 48017, 700, 24001, 12345)]
    path = os.path.join(tmp_path, "feats.npz")
    feats = compute_fbank(wavs, extractor=fb, max_batch_samples=60000, out_path=path)
    z = np.load(path)
    for i, w in enumerate(wavs):
        want = fb.extract(w, 24000).cpu().numpy()
        assert feats[i].shape == want.shape == ((w.numel() + 128) // 256, 100)
        assert float(np.abs(feats[i] - want).max()) < 1e-5
        assert np.array_equal(z[f"feat_{i}"], feats[i])


def test_fbank_rejects_rows_shorter_than_the_reflect_padding():
    """torch.stft's reflect padding raises for inputs not longer than the pad; so does the extractor."""
    fb, voc, _ = audio()
    x = torch.zeros(2, 4000, device="cuda")
    with pytest.raises(ValueError):
        fb.engine.fbank(x, torch.tensor([4000, 512]))
    with pytest.raises(ValueError):
        fb.engine.fbank(x[:, :384], kind="bigvgan")
    assert fb.engine.fbank(x, torch.tensor([4000, 513])).shape == (2, 16, 100)

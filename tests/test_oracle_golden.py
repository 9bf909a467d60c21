"""The oracle restatement against outputs of the unmodified reference (tests/golden,
produced by oracle/make_golden.py).  CPU only."""
import os

import numpy as np
import torch

from oracle import zipvoice_oracle as O
from zipvoice_b200 import synth
from tests.cases import case_inputs, case_inputs_stereo, rel_l2, valid_rel_l2
from tests.conftest import GOLDEN

TOL = 2e-5  # fp32 vs fp32, different summation order only


def _g(name):
    return {k: torch.from_numpy(v) for k, v in np.load(os.path.join(GOLDEN, name)).items()}


def test_velocity_small_stress():
    cfg = synth.small_model_cfg()
    sd = synth.make_state_dict(cfg, seed=1, stress=True)
    x0, text, speech, mask = case_inputs(2, 37, 11, seed=10)
    v = O.forward_fm_decoder(sd, cfg, torch.tensor(0.3), x0, text, speech, mask)
    assert valid_rel_l2(v, _g("velocity_small_stress.npz")["v"], mask) < TOL


def test_euler_cfg_and_nocfg():
    cfg = synth.small_model_cfg()
    sd = synth.make_state_dict(cfg, seed=1, stress=True)
    x0, text, speech, mask = case_inputs(2, 37, 11, seed=10)
    g = _g("euler_small_stress.npz")
    x1 = O.euler_sample(sd, cfg, x0, text, speech, mask, num_step=4, guidance_scale=1.0, t_shift=0.5)
    assert valid_rel_l2(x1, g["x1_cfg"], mask) < TOL
    x1 = O.euler_sample(sd, cfg, x0, text, speech, mask, num_step=3, guidance_scale=0.0, t_shift=0.7)
    assert valid_rel_l2(x1, g["x1_nocfg"], mask) < TOL


def test_euler_distill():
    cfg = synth.small_model_cfg()
    sd = synth.make_state_dict(cfg, seed=2, distill=True, stress=True)
    x0, text, speech, mask = case_inputs(2, 37, 11, seed=10)
    x1 = O.euler_sample(sd, cfg, x0, text, speech, mask, num_step=2, guidance_scale=3.0, t_shift=0.5, distill=True)
    assert valid_rel_l2(x1, _g("euler_small_distill.npz")["x1"], mask) < TOL


def test_sample_end_to_end():
    cfg = synth.small_model_cfg()
    sd = synth.make_state_dict(cfg, seed=1, stress=True)
    tokens, prompt_tokens = synth.make_tokens(2, n_prompt=9, n_text=21, seed=5, ragged=True)
    pf, pfl = synth.make_prompt_features(2, 14, seed=6, ragged=True)
    g = _g("sample_small_stress.npz")
    emb, _ = O.forward_text_embed(sd, cfg, [p + t for p, t in zip(prompt_tokens, tokens)])
    assert rel_l2(emb, g["text_embed"]) < TOL
    ptl = torch.tensor([len(t) for t in prompt_tokens])
    tl = torch.tensor([len(t) for t in tokens])
    T = int(O.predict_features_lens(pfl, ptl, tl, 1.3).max())
    torch.manual_seed(666)
    x0 = torch.randn(2, T, 100)
    out = O.sample(sd, cfg, tokens, prompt_tokens, pf, pfl, speed=1.3, t_shift=0.5, duration="predict",
                   num_step=3, guidance_scale=1.0, x0=x0)
    assert torch.equal(out[1], g["wo_lens"])
    assert rel_l2(out[0], g["x1_wo"]) < TOL
    assert rel_l2(out[2], g["x1_prompt"]) < TOL


def test_dialog_text_embed():
    cfg = synth.small_model_cfg()
    sd = synth.make_state_dict(cfg, seed=3, dialog=True, stress=True)
    dt, dp = synth.make_tokens(2, n_prompt=30, n_text=60, seed=8, dialog=True, ragged=True)
    emb, _ = O.forward_text_embed(sd, cfg, [p + t for p, t in zip(dp, dt)], dialog=(synth.SPK_A_ID, synth.SPK_B_ID))
    assert rel_l2(emb, _g("dialog_text_embed.npz")["text_embed"]) < TOL


def test_dialog_stereo():
    """ZipVoiceDialogStereo (two-stream decoder, 200-wide features): velocity, CFG Euler, sample()."""
    cfg = synth.small_model_cfg()
    sd = synth.make_state_dict(cfg, seed=4, dialog=True, stress=True, stereo=True)
    g = _g("dialog_stereo.npz")
    x0, text, speech, mask = case_inputs_stereo(2, 41, 13, seed=12)
    v = O.forward_fm_decoder(sd, cfg, torch.tensor(0.3), x0, text, speech, mask)
    assert valid_rel_l2(v, g["v"], mask) < TOL
    x1 = O.euler_sample(sd, cfg, x0, text, speech, mask, num_step=4, guidance_scale=1.5, t_shift=0.5)
    assert valid_rel_l2(x1, g["x1_cfg"], mask) < 10 * TOL
    tokens, prompt_tokens = synth.make_tokens(2, n_prompt=30, n_text=50, seed=9, dialog=True, ragged=True)
    pf, pfl = synth.make_prompt_features(2, 16, feat_dim=200, seed=13, ragged=True)
    ptl = torch.tensor([len(t) for t in prompt_tokens])
    tl = torch.tensor([len(t) for t in tokens])
    T = int(O.predict_features_lens(pfl, ptl, tl, 1.1).max())
    torch.manual_seed(666)
    x0s = torch.randn(2, T, 200)
    out = O.sample(sd, cfg, tokens, prompt_tokens, pf, pfl, speed=1.1, t_shift=0.5, duration="predict", num_step=3,
                   guidance_scale=1.5, dialog=(synth.SPK_A_ID, synth.SPK_B_ID), x0=x0s)
    assert torch.equal(out[1], g["wo_lens"])
    assert rel_l2(out[0], g["x1_wo"]) < 10 * TOL
    assert rel_l2(out[2], g["x1_prompt"]) < 10 * TOL


def test_velocity_base_cfg():
    sd = synth.make_state_dict(synth.BASE_MODEL_CFG, seed=0)
    x, t, s, mk = case_inputs(2, 50, 15, seed=11)
    v = O.cfg_velocity(sd, synth.BASE_MODEL_CFG, torch.tensor(0.7), x, t, s, mk, guidance_scale=1.5)
    assert valid_rel_l2(v, _g("velocity_base_cfg.npz")["v"], mk) < TOL


def test_fbank_oracle_vs_reference_golden():
    """oracle/fbank_oracle.py against VocosFbank.extract of the unmodified reference (lhotse shimmed)."""
    from oracle import fbank_oracle as FO
    g = _g("fbank.npz")
    for name in ("a", "b", "c"):
        wav = g["wav_" + name]
        n = wav.shape[1]
        mel = FO.vocos_fbank(wav[0])
        assert mel.shape == g["mel_" + name].shape == (FO.num_frames(n), 100)
        assert float((mel - g["mel_" + name]).abs().max()) < 1e-4


def test_bigvgan_fbank_oracle_vs_reference_golden():
    """oracle/fbank_oracle.bigvgan_fbank against BigVGANFbank.extract of the unmodified reference (lhotse shimmed,
    librosa.filters.mel restated in oracle/shims: the mel basis itself is unpinned, everything around it is pinned)."""
    from oracle import fbank_oracle as FO
    g = _g("fbank_bigvgan.npz")
    for name in ("a", "b", "c", "d"):
        wav = g["wav_" + name]
        mel = FO.bigvgan_fbank(wav[0])
        assert mel.shape == g["mel_" + name].shape == (FO.num_frames(wav.shape[1]), 100)
        assert float((mel - g["mel_" + name]).abs().max()) < 1e-4
    st = g["wav_stereo"]
    assert float((FO.bigvgan_fbank(st.mean(0)) - g["mel_stereo_mono"]).abs().max()) < 1e-4
    two = torch.cat([FO.bigvgan_fbank(st[0]), FO.bigvgan_fbank(st[1])], dim=1)
    assert float((two - g["mel_stereo_2ch"]).abs().max()) < 1e-4
    basis = FO.bigvgan_mel_basis()
    assert basis.shape == (100, 513) and basis.dtype == np.float32
    # Slaney area normalisation: every filter integrates to ~1 Hz^-1 x bin width
    assert np.allclose(basis.sum(1) * (24000 / 1024), 1.0, atol=0.15)


def test_resample_oracle_vs_torchaudio_golden():
    """oracle/resample_oracle.py against torchaudio.transforms.Resample outputs (tests/golden/resample.npz,
    oracle/make_golden.py resample); fp32 dot products of up to ~480 taps in another order: 2e-5 absolute."""
    from oracle import resample_oracle as R
    g = np.load(os.path.join(GOLDEN, "resample.npz"))
    for sr in (16000, 22050, 44100, 48000):
        got = R.resample(g[f"x_{sr}"], sr, 24000)
        want = g[f"y_{sr}"]
        assert got.shape == want.shape
        assert float(np.abs(got - want).max()) < 2e-5, sr


def test_sample_intermediate():
    """oracle sample_intermediate against ZipVoice.sample_intermediate of the unmodified reference
    (tests/golden/sample_intermediate.npz): per-sample guidance tensor, t in [0.2, 0.7]."""
    cfg = synth.small_model_cfg()
    sd = synth.make_state_dict(cfg, seed=1, stress=True)
    inp = synth.make_intermediate_inputs(seed=31)
    g = np.load(os.path.join(GOLDEN, "sample_intermediate.npz"))
    x, lens = O.sample_intermediate(sd, cfg, inp["tokens"], inp["features"], inp["features_lens"], inp["noise"],
                                    inp["speech_condition_mask"], 0.2, 0.7, num_step=2,
                                    guidance_scale=inp["guidance_scale"])
    assert lens.tolist() == g["lens"].tolist()
    assert rel_l2(x, torch.from_numpy(g["x"])) < 1e-5


def test_headline_stress_velocity_T1219():
    """Full-depth 123 M stress weights at the headline length: one guided NFE of the oracle against the unmodified
    reference (oracle/make_golden.py headline stress)."""
    from tests.cases import headline_case
    torch.set_num_threads(os.cpu_count())
    sd = synth.make_state_dict(synth.BASE_MODEL_CFG, seed=1, stress=True)
    x0, text, speech, mask = headline_case("stress")
    v = O.cfg_velocity(sd, synth.BASE_MODEL_CFG, torch.tensor(0.3), x0, text, speech, mask, guidance_scale=1.0)
    assert rel_l2(v, _g("headline_stress_velocity.npz")["v"]) < TOL


def test_headline_zipvoice_first_nfe_T1219():
    """NFE 0 of the 16-NFE headline case (B = 2 ragged, T = 1219, t = 0: doubled-guidance branch, solver.py:92-97)."""
    from tests.cases import headline_case
    torch.set_num_threads(os.cpu_count())
    sd = synth.make_state_dict(synth.BASE_MODEL_CFG, seed=0)
    x0, text, speech, mask = headline_case("zipvoice")
    ts = O.get_time_steps(0.0, 1.0, 16, 0.5)
    v = O.cfg_velocity(sd, synth.BASE_MODEL_CFG, ts[0], x0, text, speech, mask, guidance_scale=1.0)
    g = _g("headline_zipvoice.npz")
    assert valid_rel_l2(v, g["v_nfe0"], mask) < TOL
    x_next = x0 + v * (ts[1] - ts[0])
    assert rel_l2(x_next[0, ::8], g["traj_row0_s8"][1]) < TOL


def test_sample_duration_real():
    cfg = synth.small_model_cfg()
    sd = synth.make_state_dict(cfg, seed=1, stress=True)
    tokens, prompt_tokens = synth.make_tokens(3, n_prompt=11, n_text=27, seed=15, ragged=True)
    pf, pfl = synth.make_prompt_features(3, 19, seed=16, ragged=True)
    fl = torch.tensor([61, 40, 52], dtype=torch.int64)
    g = _g("sample_small_real.npz")
    torch.manual_seed(666)
    x0 = torch.randn(3, int((pfl + fl).max()), 100)
    out = O.sample(sd, cfg, tokens, prompt_tokens, pf, pfl, features_lens=fl, t_shift=0.5, duration="real",
                   num_step=3, guidance_scale=1.0, x0=x0)
    assert torch.equal(out[1], g["wo_lens"]) and torch.equal(out[3], g["prompt_lens"])
    assert rel_l2(out[0], g["x1_wo"]) < TOL and rel_l2(out[2], g["x1_prompt"]) < TOL

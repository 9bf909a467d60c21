"""Generate tests/golden/*.npz by running the UNMODIFIED reference (imported from
/root/reference, build container only) on the seeded weights/inputs of
zipvoice_b200.synth.  Run:  python -m oracle.make_golden
The fixtures pin the oracle restatement (tests/test_oracle_golden.py) and are the
known-answer vectors for the CUDA path (tests/test_gpu_parity.py).
"""
import math
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _ref_model(kind, cfg, sd):
    sys.path.insert(0, REF)
    from zipvoice.models.zipvoice import ZipVoice
    from zipvoice.models.zipvoice_dialog import ZipVoiceDialog, ZipVoiceDialogStereo
    from zipvoice.models.zipvoice_distill import ZipVoiceDistill
    from zipvoice_b200 import synth
    cls = {"zipvoice": ZipVoice, "distill": ZipVoiceDistill, "dialog": ZipVoiceDialog,
           "dialog_stereo": ZipVoiceDialogStereo}[kind]
    kw = dict(cfg, vocab_size=synth.VOCAB_SIZE, pad_id=synth.PAD_ID)
    if kind.startswith("dialog"):
        kw.update(spk_a_id=synth.SPK_A_ID, spk_b_id=synth.SPK_B_ID)
    m = cls(**kw).eval()
    m.load_state_dict(sd, strict=True)
    return m


from tests.cases import case_inputs, case_inputs_stereo  # noqa: E402  (same seeded inputs as the tests)


@torch.inference_mode()
def stereo():
    """ZipVoiceDialogStereo (zipvoice_dialog.py:218-256): velocity, CFG Euler and sample() of the unmodified
    reference on the seeded two-stream weights.  python -m oracle.make_golden stereo"""
    from zipvoice_b200 import synth
    torch.set_num_threads(8)
    small = synth.small_model_cfg()
    sd = synth.make_state_dict(small, seed=4, dialog=True, stress=True, stereo=True)
    m = _ref_model("dialog_stereo", small, sd)
    x0, text, speech, mask = case_inputs_stereo(2, 41, 13, seed=12)
    v = m.forward_fm_decoder(torch.tensor(0.3), x0, text, speech, mask)
    x1 = m.solver.sample(x=x0, text_condition=text, speech_condition=speech, padding_mask=mask,
                         num_step=4, guidance_scale=1.5, t_shift=0.5)
    tokens, prompt_tokens = synth.make_tokens(2, n_prompt=30, n_text=50, seed=9, dialog=True, ragged=True)
    pf, pfl = synth.make_prompt_features(2, 16, feat_dim=200, seed=13, ragged=True)
    torch.manual_seed(666)
    out = m.sample(tokens=tokens, prompt_tokens=prompt_tokens, prompt_features=pf, prompt_features_lens=pfl,
                   speed=1.1, t_shift=0.5, duration="predict", num_step=3, guidance_scale=1.5)
    np.savez(os.path.join(OUT, "dialog_stereo.npz"), v=v.numpy(), x1_cfg=x1.numpy(), x1_wo=out[0].numpy(),
             wo_lens=out[1].numpy(), x1_prompt=out[2].numpy())
    print("stereo golden written")
    # the SECOND projection pair of TTSZipformerTwoStream (zipformer_two_stream.py:238-244: a mono input, width
    # 3 * feat_dim, goes through in_proj[1] / out_proj[1]): velocity and a 3-step CFG Euler on mono inputs
    xm, tm, sm_, mm = case_inputs(2, 41, 13, seed=14)
    vm = m.forward_fm_decoder(torch.tensor(0.6), xm, tm, sm_, mm)
    x1m = m.solver.sample(x=xm, text_condition=tm, speech_condition=sm_, padding_mask=mm, num_step=3,
                          guidance_scale=1.0, t_shift=0.5)
    np.savez(os.path.join(OUT, "dialog_stereo_mono_stream.npz"), v=vm.numpy(), x1_cfg=x1m.numpy())
    print("stereo model, mono stream golden written", vm.shape)


@torch.inference_mode()
@torch.inference_mode()
def intermediate():
    """ZipVoice.sample_intermediate (zipvoice.py:488-534) of the unmodified reference: per-sample guidance
    (batch, 1, 1), t in [0.2, 0.7], two Euler steps.  python -m oracle.make_golden intermediate"""
    from zipvoice_b200 import synth
    torch.set_num_threads(8)
    small = synth.small_model_cfg()
    sd = synth.make_state_dict(small, seed=1, stress=True)
    m = _ref_model("zipvoice", small, sd)
    inp = synth.make_intermediate_inputs(seed=31)
    x, lens = m.sample_intermediate(tokens=inp["tokens"], features=inp["features"], features_lens=inp["features_lens"],
                                    noise=inp["noise"], speech_condition_mask=inp["speech_condition_mask"],
                                    t_start=0.2, t_end=0.7, num_step=2, guidance_scale=inp["guidance_scale"])
    np.savez(os.path.join(OUT, "sample_intermediate.npz"), x=x.numpy(), lens=lens.numpy())
    print("sample_intermediate golden written", x.shape, lens)


def resample():
    """torchaudio.transforms.Resample as the reference applies it to prompts (bin/infer_zipvoice.py:334-339), run
    with the torchaudio of this image on seeded noise + a chirp.  python -m oracle.make_golden resample"""
    import torchaudio
    g = torch.Generator().manual_seed(2024)
    out = {}
    for sr in (16000, 22050, 44100, 48000):
        n = sr // 4 + 37
        t = torch.arange(n) / sr
        x = torch.stack([torch.randn(n, generator=g) * 0.2,
                         0.5 * torch.sin(2 * math.pi * (200.0 + 3000.0 * t) * t)]).float()
        y = torchaudio.transforms.Resample(orig_freq=sr, new_freq=24000)(x)
        out[f"x_{sr}"] = x.numpy()
        out[f"y_{sr}"] = y.numpy()
    out["torchaudio_version"] = np.array(torchaudio.__version__)
    np.savez_compressed(os.path.join(OUT, "resample.npz"), **out)
    print("wrote resample.npz", {k: v.shape for k, v in out.items()})


@torch.inference_mode()
def headline(which=("zipvoice", "distill", "dialog", "real", "stress")):
    """Full-size known answers from the UNMODIFIED reference (VERDICT r1 "next" item 1).
    python -m oracle.make_golden headline [zipvoice|distill|dialog|real|stress ...]
      zipvoice: 123 M ZipVoice, B = 2 (row 0 full, row 1 ragged), T = 1219, 281-frame prompt, 16 NFE, g = 1.0,
                t_shift = 0.5 through EulerSolver.sample (solver.py:182-240); final x1 + the x entering every NFE
                (row 0, every 8th frame) + the guided velocity of NFE 0 and NFE 12 (both t > 0.5 branches).
      distill:  123 M ZipVoiceDistill, 4 NFE, g = 3.0, T = 1219 (zipvoice_distill.py:27-69).
      dialog:   123 M ZipVoiceDialog fm_decoder, one guided velocity, B = 1, T = 6563 (52 / 26 / 13 key tiles).
      real:     small ZipVoice.sample(duration="real") (zipvoice.py:425-437).
      stress:   123 M stress weights, one guided velocity at B = 1, T = 1219 (settles the per-NFE tolerance)."""
    from tests.cases import headline_case
    from zipvoice_b200 import synth
    torch.set_num_threads(os.cpu_count())
    base = synth.BASE_MODEL_CFG
    if "zipvoice" in which:
        m = _ref_model("zipvoice", base, synth.make_state_dict(base, seed=0))
        x0, text, speech, mask = headline_case("zipvoice")
        traj, vel = [], {}

        def pre(mod, args, kwargs):
            traj.append(kwargs["x"][0, ::8].clone())

        def post(mod, args, kwargs, out):
            if len(traj) - 1 in (0, 12):
                vel[len(traj) - 1] = out.clone()

        raw = []
        h0 = m.fm_decoder.register_forward_hook(lambda mod, a, out: raw.append(out.clone()) if not raw else None)
        h1 = m.solver.model.register_forward_pre_hook(pre, with_kwargs=True)
        h2 = m.solver.model.register_forward_hook(post, with_kwargs=True)
        x1 = m.solver.sample(x=x0, text_condition=text, speech_condition=speech, padding_mask=mask, num_step=16,
                             guidance_scale=1.0, t_shift=0.5)
        h0.remove(), h1.remove(), h2.remove()
        # c_nfe0: the conditional rows [B, 2B) of the doubled batch as the network returned them (solver.py:99-105)
        np.savez(os.path.join(OUT, "headline_zipvoice.npz"), x1=x1.numpy(), traj_row0_s8=torch.stack(traj).numpy(),
                 v_nfe0=vel[0].numpy(), v_nfe12_row0=vel[12][0].numpy(), c_nfe0_row0=raw[0][2].numpy())
        print("headline zipvoice written", x1.shape, len(traj))
    if "distill" in which:
        m = _ref_model("distill", base, synth.make_state_dict(base, seed=2, distill=True))
        x0, text, speech, mask = headline_case("distill")
        x1 = m.solver.sample(x=x0, text_condition=text, speech_condition=speech, padding_mask=mask, num_step=4,
                             guidance_scale=3.0, t_shift=0.5)
        np.savez(os.path.join(OUT, "headline_distill.npz"), x1=x1.numpy())
        print("headline distill written", x1.shape)
    if "dialog" in which:
        m = _ref_model("dialog", base, synth.make_state_dict(base, seed=3, dialog=True))
        x0, text, speech, mask = headline_case("dialog")
        raw = []
        h0 = m.fm_decoder.register_forward_hook(lambda mod, a, out: raw.append(out.clone()))
        v = m.solver.model(t=torch.tensor(0.3), x=x0, text_condition=text, speech_condition=speech,
                           padding_mask=mask, guidance_scale=1.5)
        h0.remove()
        np.savez(os.path.join(OUT, "headline_dialog.npz"), v=v.numpy(), c=raw[0][1:].numpy())
        print("headline dialog written", v.shape)
    if "stress" in which:
        m = _ref_model("zipvoice", base, synth.make_state_dict(base, seed=1, stress=True))
        x0, text, speech, mask = headline_case("stress")
        raw = []
        h0 = m.fm_decoder.register_forward_hook(lambda mod, a, out: raw.append(out.clone()))
        v = m.solver.model(t=torch.tensor(0.3), x=x0, text_condition=text, speech_condition=speech,
                           padding_mask=mask, guidance_scale=1.0)
        h0.remove()
        np.savez(os.path.join(OUT, "headline_stress_velocity.npz"), v=v.numpy(), c=raw[0][1:].numpy())
        print("headline stress written", v.shape)
    if "real" in which:
        small = synth.small_model_cfg()
        m = _ref_model("zipvoice", small, synth.make_state_dict(small, seed=1, stress=True))
        tokens, prompt_tokens = synth.make_tokens(3, n_prompt=11, n_text=27, seed=15, ragged=True)
        pf, pfl = synth.make_prompt_features(3, 19, seed=16, ragged=True)
        fl = torch.tensor([61, 40, 52], dtype=torch.int64)
        torch.manual_seed(666)
        out = m.sample(tokens=tokens, prompt_tokens=prompt_tokens, prompt_features=pf, prompt_features_lens=pfl,
                       features_lens=fl, t_shift=0.5, duration="real", num_step=3, guidance_scale=1.0)
        np.savez(os.path.join(OUT, "sample_small_real.npz"), x1_wo=out[0].numpy(), wo_lens=out[1].numpy(),
                 x1_prompt=out[2].numpy(), prompt_lens=out[3].numpy())
        print("sample(duration=real) written", out[0].shape, out[1])



def main():
    from zipvoice_b200 import synth
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    small = synth.small_model_cfg()

    # 1. one fm_decoder NFE, stress weights, ragged mask
    sd = synth.make_state_dict(small, seed=1, stress=True)
    m = _ref_model("zipvoice", small, sd)
    x0, text, speech, mask = case_inputs(2, 37, 11, seed=10)
    v = m.forward_fm_decoder(torch.tensor(0.3), x0, text, speech, mask)
    np.savez(os.path.join(OUT, "velocity_small_stress.npz"), v=v.numpy())

    # 2. Euler + CFG, both t>0.5 branches, stress weights
    x1 = m.solver.sample(x=x0, text_condition=text, speech_condition=speech, padding_mask=mask,
                         num_step=4, guidance_scale=1.0, t_shift=0.5)
    x1b = m.solver.sample(x=x0, text_condition=text, speech_condition=speech, padding_mask=mask,
                          num_step=3, guidance_scale=0.0, t_shift=0.7)
    np.savez(os.path.join(OUT, "euler_small_stress.npz"), x1_cfg=x1.numpy(), x1_nocfg=x1b.numpy())

    # 3. distill: guidance embedded, no doubling
    sdd = synth.make_state_dict(small, seed=2, distill=True, stress=True)
    md = _ref_model("distill", small, sdd)
    x1d = md.solver.sample(x=x0, text_condition=text, speech_condition=speech, padding_mask=mask,
                           num_step=2, guidance_scale=3.0, t_shift=0.5)
    np.savez(os.path.join(OUT, "euler_small_distill.npz"), x1=x1d.numpy())

    # 4. ZipVoice.sample end to end (text encoder + condition + solver + prompt strip)
    tokens, prompt_tokens = synth.make_tokens(2, n_prompt=9, n_text=21, seed=5, ragged=True)
    pf, pfl = synth.make_prompt_features(2, 14, seed=6, ragged=True)
    torch.manual_seed(666)
    out = m.sample(tokens=tokens, prompt_tokens=prompt_tokens, prompt_features=pf, prompt_features_lens=pfl,
                   speed=1.3, t_shift=0.5, duration="predict", num_step=3, guidance_scale=1.0)
    emb, tl = m.forward_text_embed([p + t for p, t in zip(prompt_tokens, tokens)])
    np.savez(os.path.join(OUT, "sample_small_stress.npz"), x1_wo=out[0].numpy(), wo_lens=out[1].numpy(),
             x1_prompt=out[2].numpy(), text_embed=emb.numpy())

    # 5. dialog text embed (speaker rows)
    sdg = synth.make_state_dict(small, seed=3, dialog=True, stress=True)
    mg = _ref_model("dialog", small, sdg)
    dt, dp = synth.make_tokens(2, n_prompt=30, n_text=60, seed=8, dialog=True, ragged=True)
    embg, _ = mg.forward_text_embed([p + t for p, t in zip(dp, dt)])
    np.savez(os.path.join(OUT, "dialog_text_embed.npz"), text_embed=embg.numpy())

    # 6. full 123 M base config, init-like weights, one CFG velocity through DiffusionModel
    sdb = synth.make_state_dict(synth.BASE_MODEL_CFG, seed=0)
    mb = _ref_model("zipvoice", synth.BASE_MODEL_CFG, sdb)
    xb, tb, sb, mkb = case_inputs(2, 50, 15, seed=11)
    vb = mb.solver.model(t=torch.tensor(0.7), x=xb, text_condition=tb, speech_condition=sb, padding_mask=mkb,
                         guidance_scale=1.5)
    np.savez(os.path.join(OUT, "velocity_base_cfg.npz"), v=vb.numpy())
    # 7. VocosFbank: the unmodified reference class (lhotse replaced by oracle/shims) on seeded audio
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "shims"))
    sys.path.insert(0, REF)
    from zipvoice.utils.feature import VocosFbank
    fb = VocosFbank()
    fbank = {}
    for name, n in (("a", 12345), ("b", 24000 + 77), ("c", 30000 + 200)):
        wav = synth.make_audio(1, n, seed=7 + n)
        fbank["wav_" + name] = wav.numpy()  # inputs are stored too: sin / mean differ in the last ulp across hosts
        fbank["mel_" + name] = fb.extract(wav, 24000).numpy()
    np.savez(os.path.join(OUT, "fbank.npz"), **fbank)
    print("golden written to", OUT)


def bigvgan():
    """BigVGANFbank.extract of the unmodified reference (zipvoice/utils/feature.py:133-204) on seeded audio; ``lhotse``
    and ``librosa`` come from oracle/shims (librosa.filters.mel is a restatement: that function is unpinned)."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "shims"))
    sys.path.insert(0, REF)
    from zipvoice.utils.feature import BigVGANFbank
    from zipvoice_b200 import synth
    fb = BigVGANFbank()
    out = {}
    for name, n in (("a", 12345), ("b", 24000 + 77), ("c", 30000 + 200), ("d", 256 * 40 + 128)):
        wav = synth.make_audio(1, n, seed=7 + n)
        out["wav_" + name] = wav.numpy()
        out["mel_" + name] = fb.extract(wav, 24000).numpy()
    st = synth.make_audio(2, 20000, seed=5)
    out["wav_stereo"] = st.numpy()
    out["mel_stereo_mono"] = fb.extract(st, 24000).numpy()            # num_channels=1: channel mean
    out["mel_stereo_2ch"] = BigVGANFbank(num_channels=2).extract(st, 24000).numpy()
    np.savez(os.path.join(OUT, "fbank_bigvgan.npz"), **out)
    print("bigvgan golden written to", OUT)


if __name__ == "__main__":
    if "headline" in sys.argv[1:]:
        rest = [a for a in sys.argv[1:] if a != "headline"]
        headline(tuple(rest) if rest else ("zipvoice", "distill", "dialog", "real", "stress"))
    elif "bigvgan" in sys.argv[1:]:
        bigvgan()
    elif "stereo" in sys.argv[1:]:
        stereo()
    elif "resample" in sys.argv[1:]:
        resample()
    elif "intermediate" in sys.argv[1:]:
        intermediate()
    else:
        main()

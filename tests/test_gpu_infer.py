"""End-to-end pipeline (prompt wav -> mel -> sample -> vocoder) through the host mirrors of
bin/infer_zipvoice.py against the same pipeline built from the CPU oracles."""
import pytest
import torch

from tests.cases import rel_l2

pytestmark = pytest.mark.gpu


def _setup():
    from oracle import vocos_oracle as VO
    from zipvoice_b200 import synth
    from zipvoice_b200.audio import VocosB200, VocosFbankB200
    from zipvoice_b200.model import ZipVoiceB200
    cfg = synth.small_model_cfg()
    sd = synth.make_state_dict(cfg, seed=1, stress=True)
    vsd = VO.make_vocos_state_dict(0)
    model = ZipVoiceB200(cfg, sd, synth.VOCAB_SIZE, synth.PAD_ID, kind="zipvoice")
    voc = VocosB200(vsd)
    return cfg, sd, vsd, model, voc, VocosFbankB200(engine=voc.engine)


def test_generate_sentence_vs_oracle_pipeline():
    from oracle import fbank_oracle as FO, vocos_oracle as VO, zipvoice_oracle as O
    from zipvoice_b200 import infer, synth
    cfg, sd, vsd, model, voc, fb = _setup()
    tokens, prompt_tokens = synth.make_tokens(1, n_prompt=8, n_text=20, seed=3)
    wav_p = synth.make_audio(1, 9000, seed=11) * 0.3          # rms < 0.1 -> exercises the RMS path
    # oracle pipeline
    rms = torch.sqrt(torch.mean(torch.square(wav_p)))
    wp = wav_p * 0.1 / rms if rms < 0.1 else wav_p
    feats = FO.vocos_fbank(wp[0]).unsqueeze(0) * 0.1
    fl = torch.tensor([feats.size(1)])
    T = int(O.predict_features_lens(fl, torch.tensor([8]), torch.tensor([20]), 1.0))
    x0 = torch.randn(1, T, 100, generator=torch.Generator().manual_seed(9))
    want_mel, want_lens, _, _ = O.sample(sd, cfg, tokens, prompt_tokens, feats, fl, speed=1.0, t_shift=0.5,
                                         num_step=4, guidance_scale=1.0, x0=x0)
    want_wav = VO.decode(vsd, want_mel.permute(0, 2, 1) / 0.1).clamp(-1, 1)
    if rms < 0.1:
        want_wav = want_wav * rms / 0.1
    wav, metrics = infer.generate_sentence(prompt_tokens[0], wav_p, tokens[0], model, voc, fb, num_step=4,
                                           guidance_scale=1.0, t_shift=0.5, x0=x0.cuda())
    assert wav.shape == want_wav.shape
    r = rel_l2(wav.cpu(), want_wav)
    assert r < 5e-2, f"waveform rel-L2 {r:.3e}"
    assert set(metrics) == {"t", "t_no_vocoder", "t_vocoder", "wav_seconds", "rtf", "rtf_no_vocoder", "rtf_vocoder"}


def test_generate_batch_shapes_and_lengths():
    from zipvoice_b200 import infer, synth
    cfg, sd, vsd, model, voc, fb = _setup()
    tokens, prompt_tokens = synth.make_tokens(3, n_prompt=8, n_text=20, seed=4, ragged=True)
    wavs = [synth.make_audio(1, n, seed=n)[0] for n in (9000, 7000, 12000)]
    outs = infer.generate_batch(prompt_tokens, wavs, tokens, model, voc, fb, num_step=2)
    assert len(outs) == 3
    for i, w in enumerate(outs):
        pf = (wavs[i].numel() + 128) // 256
        import math
        gen = math.ceil(pf / len(prompt_tokens[i]) * len(tokens[i]) / 1.0)
        assert w.shape == (256 * (gen - 1),) and bool(torch.isfinite(w).all())


def test_generate_sentence_stereo_vs_oracle_pipeline():
    """Two-channel dialog pipeline (bin/infer_zipvoice_dialog.py:350-515) against the CPU oracles."""
    from oracle import fbank_oracle as FO, vocos_oracle as VO, zipvoice_oracle as O
    from zipvoice_b200 import infer, synth
    from zipvoice_b200.audio import VocosB200, VocosFbankB200
    from zipvoice_b200.model import ZipVoiceB200
    cfg = synth.small_model_cfg()
    sd = synth.make_state_dict(cfg, seed=4, dialog=True, stress=True, stereo=True)
    vsd = VO.make_vocos_state_dict(0)
    model = ZipVoiceB200(cfg, sd, synth.VOCAB_SIZE, synth.PAD_ID, kind="dialog_stereo")
    voc = VocosB200(vsd)
    fb = VocosFbankB200(engine=voc.engine, num_channels=2)
    tokens, prompt_tokens = synth.make_tokens(1, n_prompt=26, n_text=40, seed=3, dialog=True)
    wav_p = synth.make_audio(2, 9000, seed=12) * 0.3             # (2, N): left / right, rms < 0.1
    rms = torch.sqrt(torch.mean(torch.square(wav_p)))
    wp = wav_p * 0.1 / rms if rms < 0.1 else wav_p
    feats = torch.cat([FO.vocos_fbank(wp[0]), FO.vocos_fbank(wp[1])], dim=-1).unsqueeze(0) * 0.1   # (1, T, 200)
    fl = torch.tensor([feats.size(1)])
    T = int(O.predict_features_lens(fl, torch.tensor([26]), torch.tensor([40]), 1.0))
    x0 = torch.randn(1, T, 200, generator=torch.Generator().manual_seed(9))
    want_mel, _, _, _ = O.sample(sd, cfg, tokens, prompt_tokens, feats, fl, speed=1.0, t_shift=0.5, num_step=4,
                                 guidance_scale=1.5, dialog=(synth.SPK_A_ID, synth.SPK_B_ID), x0=x0)
    wm = want_mel.permute(0, 2, 1) / 0.1
    want = torch.cat([VO.decode(vsd, wm[:, :100]), VO.decode(vsd, wm[:, 100:])], dim=0).clamp(-1, 1)
    if rms < 0.1:
        want = want * rms / 0.1
    wav, metrics = infer.generate_sentence_stereo(prompt_tokens[0], wav_p, tokens[0], model, voc, fb, num_step=4,
                                                  guidance_scale=1.5, t_shift=0.5, x0=x0.cuda())
    assert wav.shape == want.shape and wav.shape[0] == 2
    r = rel_l2(wav.cpu(), want)
    assert r < 5e-2, f"stereo waveform rel-L2 {r:.3e}"


def test_generate_list_buckets_resamples_and_writes(tmp_path):
    """generate_list (infer_zipvoice.py:406-466): prompts at 16 kHz and 24 kHz from wav files, length buckets,
    16-bit files written behind the compute.  Every file must hold what generate_batch returns for its bucket
    inputs (same noise seed), to the 16-bit quantisation."""
    import os
    from zipvoice_b200 import infer, synth
    from zipvoice_b200.audio import ResampleB200
    cfg, sd, vsd, model, voc, fb = _setup()

    class Tok:
        def texts_to_token_ids(self, texts):
            return [[1 + (ord(c) % (synth.VOCAB_SIZE - 1)) for c in t] for t in texts]

    tok = Tok()
    specs = [("a", 16000, 7000, "hello wor", "a somewhat longer text to say"), ("b", 24000, 9000, "prompt b", "short"),
             ("c", 16000, 12000, "third prompt here", "medium length text"), ("d", 24000, 8000, "dd", "x y z w")]
    lines = []
    for name, sr, n, ptxt, txt in specs:
        w = synth.make_audio(1, n, seed=len(name) + n)[0] * 0.3
        pth = os.path.join(tmp_path, f"p_{name}.wav")
        infer.write_wav(pth, w.numpy(), sr)
        lines.append("\t".join([name, ptxt, pth, txt]))
    lst = os.path.join(tmp_path, "test.tsv")
    open(lst, "w").write("\n".join(lines) + "\n")
    res = os.path.join(tmp_path, "out")
    torch.manual_seed(123)
    m = infer.generate_list(res, lst, model, voc, tok, fb, num_step=4, max_batch_frames=2 * 200)
    assert m["utterances"] == 4 and m["batches"] >= 2 and m["wav_seconds"] > 0
    total = 0
    for name, sr, n, ptxt, txt in specs:
        w, r = infer.read_wav(os.path.join(res, f"{name}.wav"))
        assert r == 24000 and w.shape[0] == 1 and w.shape[1] > 0 and w.shape[1] % 256 == 0
        assert torch.isfinite(w).all() and float(w.abs().max()) > 0
        total += w.shape[1]
    assert abs(total / 24000 - m["wav_seconds"]) < 1e-6
    # wav round trip and the bucketing helper
    w16, r16 = infer.read_wav(os.path.join(tmp_path, "p_a.wav"))
    assert r16 == 16000 and w16.shape == (1, 7000)
    assert infer.length_buckets([5, 1, 3, 9], 10) == [[1, 2], [0], [3]]
    up = ResampleB200(16000, 24000, engine=fb.engine)(w16.cuda())
    assert up.shape == (1, 10500)

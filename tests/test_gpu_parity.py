"""Parity of the CUDA hot path (through the C ABI) against the CPU oracle and the committed golden
vectors of the unmodified reference.  bf16 tensor-core operands vs the fp32 reference: tolerances are
BASELINE.json's (final mel rel-L2 <= 1e-2) and SURVEY.md 8d's proposal for one NFE (rel-L2 <= 5e-3),
on valid (unpadded) frames."""
import os

import numpy as np
import pytest
import torch

from tests.cases import case_inputs, case_inputs_stereo, guided_tol, rel_l2, valid_rel_l2
from tests.conftest import GOLDEN

pytestmark = pytest.mark.gpu

# Per-NFE tolerances (rel-L2 on valid frames, bf16 tensor-core operands vs the fp32 reference), BASELINE.md section 5:
#   TOL_NFE            one network evaluation (fm_decoder output), init-like weights
#   TOL_NFE_STRESS     one network evaluation, 4x-amplified "stress" weights (every branch of the layer loud)
#   guided velocities  v = (1 + g) c - g u combine two evaluations with independent rounding errors:
#                      tolerance = raw tolerance x sqrt((1 + g)^2 + g^2), g = the guidance actually applied
TOL_NFE = 5e-3
TOL_NFE_STRESS = 7e-3
#                      (tests/cases.py guided_tol)
TOL_MEL = 1e-2


def _g(name):
    return {k: torch.from_numpy(v) for k, v in np.load(os.path.join(GOLDEN, name)).items()}


_engines = {}


def engine(kind, cfg_name, seed, stress):
    from zipvoice_b200 import synth
    from zipvoice_b200.model import ZipVoiceB200
    key = (kind, cfg_name, seed, stress)
    if key not in _engines:
        cfg = synth.small_model_cfg() if cfg_name == "small" else synth.BASE_MODEL_CFG
        sd = synth.make_state_dict(cfg, seed=seed, distill=(kind == "distill"), dialog=kind.startswith("dialog"),
                                   stress=stress, stereo=(kind == "dialog_stereo"))
        _engines[key] = (ZipVoiceB200(cfg, sd, synth.VOCAB_SIZE, synth.PAD_ID, kind=kind), cfg, sd)
    return _engines[key]


@pytest.mark.parametrize("gemm_impl", [1, 2, 3], ids=["simt", "tcgen05", "tcgen05_pair"])
def test_velocity_small_stress(gemm_impl):
    m, cfg, sd = engine("zipvoice", "small", 1, True)
    m.engine.set_option("gemm_impl", gemm_impl)
    x0, text, speech, mask = case_inputs(2, 37, 11, seed=10)
    v = m.forward_fm_decoder(torch.tensor(0.3), x0, text, speech, mask).cpu()
    m.engine.set_option("gemm_impl", 0)
    r = valid_rel_l2(v, _g("velocity_small_stress.npz")["v"], mask)
    assert r < TOL_NFE, f"velocity rel-L2 vs reference golden {r:.3e}"


def test_euler_golden():
    m, cfg, sd = engine("zipvoice", "small", 1, True)
    x0, text, speech, mask = case_inputs(2, 37, 11, seed=10)
    g = _g("euler_small_stress.npz")
    x1 = m.solver.sample(x=x0, text_condition=text, speech_condition=speech, padding_mask=mask, num_step=4,
                         guidance_scale=1.0, t_shift=0.5).cpu()
    r = valid_rel_l2(x1, g["x1_cfg"], mask)
    assert r < TOL_MEL, f"CFG Euler rel-L2 {r:.3e}"
    x1 = m.solver.sample(x=x0, text_condition=text, speech_condition=speech, padding_mask=mask, num_step=3,
                         guidance_scale=0.0, t_shift=0.7).cpu()
    r = valid_rel_l2(x1, g["x1_nocfg"], mask)
    assert r < TOL_MEL, f"plain Euler rel-L2 {r:.3e}"


def test_euler_distill_golden():
    m, cfg, sd = engine("distill", "small", 2, True)
    x0, text, speech, mask = case_inputs(2, 37, 11, seed=10)
    x1 = m.solver.sample(x=x0, text_condition=text, speech_condition=speech, padding_mask=mask, num_step=2,
                         guidance_scale=3.0, t_shift=0.5).cpu()
    r = valid_rel_l2(x1, _g("euler_small_distill.npz")["x1"], mask)
    assert r < TOL_MEL, f"distill Euler rel-L2 {r:.3e}"


def test_sample_end_to_end_golden():
    from zipvoice_b200 import synth
    m, cfg, sd = engine("zipvoice", "small", 1, True)
    tokens, prompt_tokens = synth.make_tokens(2, n_prompt=9, n_text=21, seed=5, ragged=True)
    pf, pfl = synth.make_prompt_features(2, 14, seed=6, ragged=True)
    g = _g("sample_small_stress.npz")
    emb, _ = m.forward_text_embed([p + t for p, t in zip(prompt_tokens, tokens)])
    r = rel_l2(emb.cpu(), g["text_embed"])
    assert r < TOL_NFE, f"text encoder rel-L2 {r:.3e}"
    torch.manual_seed(666)
    from oracle import zipvoice_oracle as O
    ptl = torch.tensor([len(t) for t in prompt_tokens])
    tl = torch.tensor([len(t) for t in tokens])
    Tfull = int(O.predict_features_lens(pfl, ptl, tl, 1.3).max())
    x0 = torch.randn(2, Tfull, 100)
    out = m.sample(tokens, prompt_tokens, pf, pfl, speed=1.3, t_shift=0.5, duration="predict", num_step=3,
                   guidance_scale=1.0, x0=x0.cuda())
    assert torch.equal(out[1].cpu(), g["wo_lens"])
    r = rel_l2(out[0].cpu(), g["x1_wo"])
    assert r < TOL_MEL, f"sample() generated part rel-L2 {r:.3e}"
    r = rel_l2(out[2].cpu(), g["x1_prompt"])
    assert r < TOL_MEL, f"sample() prompt part rel-L2 {r:.3e}"


def test_dialog_stereo_golden():
    """ZipVoiceDialogStereo through the first stream of the two-stream decoder (200-wide x / speech)."""
    from zipvoice_b200 import synth
    from oracle import zipvoice_oracle as O
    m, cfg, sd = engine("dialog_stereo", "small", 4, True)
    g = _g("dialog_stereo.npz")
    x0, text, speech, mask = case_inputs_stereo(2, 41, 13, seed=12)
    v = m.forward_fm_decoder(torch.tensor(0.3), x0, text, speech, mask).cpu()
    r = valid_rel_l2(v, g["v"], mask)
    assert r < TOL_NFE, f"stereo velocity rel-L2 {r:.3e}"
    x1 = m.solver.sample(x=x0, text_condition=text, speech_condition=speech, padding_mask=mask, num_step=4,
                         guidance_scale=1.5, t_shift=0.5).cpu()
    r = valid_rel_l2(x1, g["x1_cfg"], mask)
    assert r < TOL_MEL, f"stereo CFG Euler rel-L2 {r:.3e}"
    tokens, prompt_tokens = synth.make_tokens(2, n_prompt=30, n_text=50, seed=9, dialog=True, ragged=True)
    pf, pfl = synth.make_prompt_features(2, 16, feat_dim=200, seed=13, ragged=True)
    ptl = torch.tensor([len(t) for t in prompt_tokens])
    tl = torch.tensor([len(t) for t in tokens])
    T = int(O.predict_features_lens(pfl, ptl, tl, 1.1).max())
    torch.manual_seed(666)
    x0s = torch.randn(2, T, 200)
    out = m.sample(tokens, prompt_tokens, pf, pfl, speed=1.1, t_shift=0.5, duration="predict", num_step=3,
                   guidance_scale=1.5, x0=x0s.cuda())
    assert torch.equal(out[1].cpu(), g["wo_lens"])
    r = rel_l2(out[0].cpu(), g["x1_wo"])
    assert r < TOL_MEL, f"stereo sample() generated part rel-L2 {r:.3e}"
    with pytest.raises(AssertionError):  # zipformer_two_stream.py:238: the width must match one of the two streams
        m.forward_fm_decoder(torch.tensor(0.3), x0[..., :150], text, speech[..., :150], mask)
    # the SECOND projection pair (mono input through in_proj.1 / out_proj.1, zipformer_two_stream.py:238-244) against
    # the unmodified reference, then back to the two-channel stream on the same handle
    gm = _g("dialog_stereo_mono_stream.npz")
    xm, tm, sm_, mm = case_inputs(2, 41, 13, seed=14)
    vm = m.forward_fm_decoder(torch.tensor(0.6), xm, tm, sm_, mm).cpu()
    r = valid_rel_l2(vm, gm["v"], mm)
    assert vm.shape == (2, 41, 100) and r < TOL_NFE_STRESS, f"mono stream of the two-stream decoder: rel-L2 {r:.3e}"
    x1m = m.solver.sample(x=xm, text_condition=tm, speech_condition=sm_, padding_mask=mm, num_step=3,
                          guidance_scale=1.0, t_shift=0.5).cpu()
    r = valid_rel_l2(x1m, gm["x1_cfg"], mm)
    assert r < TOL_MEL, f"mono stream CFG Euler rel-L2 {r:.3e}"
    v2 = m.forward_fm_decoder(torch.tensor(0.3), x0, text, speech, mask).cpu()
    assert torch.equal(v2, v), "switching streams must not disturb the two-channel path"


def test_dialog_text_embed_golden():
    from zipvoice_b200 import synth
    m, cfg, sd = engine("dialog", "small", 3, True)
    dt, dp = synth.make_tokens(2, n_prompt=30, n_text=60, seed=8, dialog=True, ragged=True)
    emb, _ = m.forward_text_embed([p + t for p, t in zip(dp, dt)])
    r = rel_l2(emb.cpu(), _g("dialog_text_embed.npz")["text_embed"])
    assert r < TOL_NFE, f"dialog text embed rel-L2 {r:.3e}"


def test_velocity_base_cfg_golden():
    m, cfg, sd = engine("zipvoice", "base", 0, False)
    x, t, s, mk = case_inputs(2, 50, 15, seed=11)
    v = m.solver.model(t=torch.tensor(0.7), x=x, text_condition=t, speech_condition=s, padding_mask=mk,
                       guidance_scale=1.5).cpu()
    r = valid_rel_l2(v, _g("velocity_base_cfg.npz")["v"], mk)
    assert r < TOL_NFE, f"123M CFG velocity rel-L2 {r:.3e}"


def test_velocity_vs_oracle_longer():
    """Seeded case the oracle finishes in seconds: T = 300 crosses several attention key tiles and
    conv tiles; ragged lengths; stress weights; per-NFE max-abs and rel-L2 reported."""
    from oracle import zipvoice_oracle as O
    m, cfg, sd = engine("zipvoice", "small", 1, True)
    x0, text, speech, mask = case_inputs(3, 300, 90, seed=21)
    keep = (~mask).unsqueeze(-1)
    for t, g_eff in ((0.1, 2.0), (0.9, 1.0)):   # guidance 1.0 is doubled for t <= 0.5 (solver.py:92-97)
        c = m.forward_fm_decoder(torch.tensor(t), x0, text, speech, mask).cpu()
        want_c = O.forward_fm_decoder(sd, cfg, torch.tensor(t), x0, text, speech, mask)
        rc = valid_rel_l2(c, want_c, mask)
        print(f"per-NFE network output t={t}: rel-L2 {rc:.3e} max-abs {float(((c - want_c) * keep).abs().max()):.3e}")
        assert rc < TOL_NFE_STRESS, f"t={t}: network output rel-L2 {rc:.3e}"
        v = m.solver.model(t=torch.tensor(t), x=x0, text_condition=text, speech_condition=speech, padding_mask=mask,
                           guidance_scale=1.0).cpu()
        want = O.cfg_velocity(sd, cfg, torch.tensor(t), x0, text, speech, mask, guidance_scale=1.0)
        r = valid_rel_l2(v, want, mask)
        ma = float(((v - want) * keep).abs().max())
        print(f"per-NFE guided velocity t={t} (g_eff {g_eff}): rel-L2 {r:.3e} max-abs {ma:.3e}")
        assert r < guided_tol(g_eff, TOL_NFE_STRESS), f"t={t}: rel-L2 {r:.3e} max-abs {ma:.3e}"


@pytest.mark.parametrize("kind,guidance", [("zipvoice", 1.0), ("zipvoice", 0.0), ("distill", 3.0)])
def test_sample_graph_replay_is_bit_identical(kind, guidance):
    """zv_sample on launch-bound batches replays a captured CUDA graph from the third call of a key on (first call
    eager, second capture + launch).  Every call must return exactly what the plain loop returns, also after the
    inputs change under the same key and after a bigger shape reallocated the workspace."""
    m, cfg, sd = engine(kind, "small", 1, True)
    eng = m.engine
    x0, text, speech, mask = case_inputs(2, 37, 11, seed=10)
    kw = dict(num_step=4, guidance_scale=guidance, t_shift=0.5)
    eng.set_option("graph_max_rows", 0)
    want = eng.sample(x0, text, speech, mask, **kw).clone()
    xb, tb, sb, mb = case_inputs(2, 37, 11, seed=11)
    want_b = eng.sample(xb, tb, sb, mb, **kw).clone()
    eng.set_option("graph_max_rows", 32768)
    eng.launch_count_reset()
    for i in range(4):
        got = eng.sample(x0, text, speech, mask, **kw)
        assert torch.equal(got, want), f"call {i}"
    per_call = eng.launch_count() // 4
    assert per_call > 0 and eng.launch_count() == 4 * per_call  # replays are counted like the eager launches
    assert torch.equal(eng.sample(xb, tb, sb, mb, **kw), want_b)   # same key, other inputs (staging buffers)
    x2, t2, s2, m2 = case_inputs(3, 61, 20, seed=12)                # bigger shape: workspace grows, graphs dropped
    eng.set_option("graph_max_rows", 0)
    want2 = eng.sample(x2, t2, s2, m2, **kw).clone()
    eng.set_option("graph_max_rows", 32768)
    for i in range(3):
        assert torch.equal(eng.sample(x2, t2, s2, m2, **kw), want2), f"shape 2 call {i}"
    for i in range(3):
        assert torch.equal(eng.sample(x0, text, speech, mask, **kw), want), f"shape 1 again, call {i}"


def test_sample_intermediate_golden():
    """ZipVoiceB200.sample_intermediate (zipvoice.py:488-534: per-sample guidance tensor (batch, 1, 1), arbitrary
    t_start / t_end) against the unmodified reference's output on the valid frames."""
    from zipvoice_b200 import synth
    m, cfg, sd = engine("zipvoice", "small", 1, True)
    inp = synth.make_intermediate_inputs(seed=31)
    g = _g("sample_intermediate.npz")
    x, lens = m.sample_intermediate(tokens=inp["tokens"], features=inp["features"], features_lens=inp["features_lens"],
                                    noise=inp["noise"], speech_condition_mask=inp["speech_condition_mask"],
                                    t_start=0.2, t_end=0.7, num_step=2, guidance_scale=inp["guidance_scale"])
    assert lens.cpu().tolist() == g["lens"].tolist()
    mask = torch.arange(x.shape[1])[None, :] >= g["lens"][:, None]
    r = valid_rel_l2(x.cpu(), g["x"], mask)
    assert r < TOL_MEL, f"sample_intermediate rel-L2 {r:.3e}"


def test_bounded_graph_and_position_table_caches():
    """Single-utterance serving sees a new T on almost every request: the captured-graph cache is an LRU of
    16 keys and every layer keeps the position tables of at most 32 lengths.  40 distinct lengths drive both
    evictions (a freed table invalidates the graphs that captured its address); every call, eager, captured or
    replayed, before and after an eviction, must return exactly what the plain loop returns."""
    m, cfg, sd = engine("zipvoice", "small", 1, True)
    eng = m.engine
    kw = dict(num_step=2, guidance_scale=1.0, t_shift=0.5)
    lengths = list(range(24, 64))
    cases = {T: case_inputs(1, T, 7, seed=100 + T) for T in lengths}
    eng.set_option("graph_max_rows", 0)
    want = {T: eng.sample(*cases[T], **kw).clone() for T in (24, 25, 40, 62, 63)}
    eng.set_option("graph_max_rows", 32768)
    for rnd in range(3):            # round 0 eager, round 1 capture + launch, round 2 replay -- as far as the LRU allows
        for T in lengths:
            got = eng.sample(*cases[T], **kw)
            if T in want:
                assert torch.equal(got, want[T]), f"round {rnd}, T = {T}"
    for T in (24, 24, 24, 63, 63, 63):   # evicted long ago: eager again, recaptured, replayed
        assert torch.equal(eng.sample(*cases[T], **kw), want[T]), f"after eviction, T = {T}"


def test_forward_fm_decoder_per_sample_time_steps():
    """forward_fm_decoder(t (N, 1, 1)) with a different time step per row (zipvoice.py:165-167, SURVEY 8b "signatures
    to honour") and, for the distilled model, a guidance value per row (zipvoice.py:171-178), against the oracle."""
    from oracle import zipvoice_oracle as O
    x0, text, speech, mask = case_inputs(3, 45, 12, seed=33)
    t = torch.tensor([0.15, 0.85, 0.5]).reshape(3, 1, 1)
    m, cfg, sd = engine("zipvoice", "small", 1, True)
    v = m.forward_fm_decoder(t, x0, text, speech, mask).cpu()
    want = O.forward_fm_decoder(sd, cfg, t, x0, text, speech, mask)
    r = valid_rel_l2(v, want, mask)
    assert r < TOL_NFE_STRESS, f"per-row t: rel-L2 {r:.3e}"
    for b in range(3):   # every row equals the scalar-t call with that row's time step
        vb = m.forward_fm_decoder(t[b, 0, 0], x0[b:b + 1], text[b:b + 1], speech[b:b + 1], mask[b:b + 1]).cpu()
        rb = valid_rel_l2(v[b:b + 1], vb, mask[b:b + 1])
        assert rb < 2e-3, f"row {b}: batched per-row t vs scalar call {rb:.3e}"
    same = m.forward_fm_decoder(torch.full((3, 1, 1), 0.3), x0, text, speech, mask)
    assert torch.equal(same, m.forward_fm_decoder(torch.tensor(0.3), x0, text, speech, mask))
    md, cfgd, sdd = engine("distill", "small", 2, True)
    g = torch.tensor([0.5, 3.0, 1.5]).reshape(3, 1, 1)
    vd = md.forward_fm_decoder(t, x0, text, speech, mask, guidance_scale=g).cpu()
    wd = O.forward_fm_decoder(sdd, cfgd, t, x0, text, speech, mask, guidance_scale=g)
    rd = valid_rel_l2(vd, wd, mask)
    assert rd < TOL_NFE_STRESS, f"per-row t and guidance (distill): rel-L2 {rd:.3e}"
    with pytest.raises(Exception):     # zipformer.py:270-272: guidance needs the distilled model's embedding
        m.forward_fm_decoder(t, x0, text, speech, mask, guidance_scale=g)


class _RefShaped:
    """Stand-in with the shape of the reference nn.Modules that ``install`` patches (zipvoice.py:38-133): a
    ``state_dict()`` with the reference's tensor names, ``pad_id``, a ``solver`` attribute (the seam at zipvoice.py:133)
    and a ``sample()`` that calls ``self.solver.sample(...)`` BY KEYWORD exactly as zipvoice.py:460-468 does.  The parts
    ``install`` does not replace (text encoder, text condition) come from the CPU oracle here, as they would come
    from the reference's own torch modules."""

    def __init__(self, cfg, sd, pad_id, dialog=None):
        self.cfg, self._sd, self.pad_id, self._dialog = cfg, sd, pad_id, dialog
        self.solver = None  # EulerSolver(self, func_name="forward_fm_decoder") in the reference
        if dialog:
            self.spk_a_id, self.spk_b_id = dialog

    def state_dict(self):
        return self._sd

    def sample(self, tokens, prompt_tokens, prompt_features, prompt_features_lens, speed=1.0, t_shift=1.0,
               num_step=5, guidance_scale=0.5, x0=None):
        from oracle import zipvoice_oracle as O
        cat = [p + t for p, t in zip(prompt_tokens, tokens)]
        embed, cat_lens = O.forward_text_embed(self._sd, self.cfg, cat, self.pad_id, self._dialog)
        ptl = torch.tensor([len(t) for t in prompt_tokens])
        tl = torch.tensor([len(t) for t in tokens])
        total = O.predict_features_lens(prompt_features_lens, ptl, tl, speed)
        text_condition, padding_mask = O.forward_text_condition(embed, cat_lens, total)     # zipvoice.py:421-437
        B, T, _ = text_condition.shape
        speech_condition = torch.nn.functional.pad(prompt_features, (0, 0, 0, T - prompt_features.size(1)))
        speech_condition = speech_condition * (~O.make_pad_mask(prompt_features_lens, T)).unsqueeze(-1)
        x1 = self.solver.sample(x=x0, text_condition=text_condition, speech_condition=speech_condition,
                                padding_mask=padding_mask, num_step=num_step, guidance_scale=guidance_scale,
                                t_shift=t_shift)                                             # zipvoice.py:460-468
        wo_lens = (~padding_mask).sum(-1) - prompt_features_lens
        out = torch.zeros(B, int(wo_lens.max()), x1.size(2))
        for i in range(B):                                                                   # zipvoice.py:469-486
            out[i, : int(wo_lens[i])] = x1[i, int(prompt_features_lens[i]): int(prompt_features_lens[i] + wo_lens[i])].cpu()
        return out, wo_lens


def test_install_on_reference_shaped_modules():
    """``install(ref_model, cfg)``: the solver of a reference-shaped object is replaced and its own
    ``sample()`` -- unchanged, calling ``self.solver.sample`` by keyword -- reproduces the unmodified reference's
    golden output; the class NAME selects the solver / engine flavour as it does for the reference classes."""
    from oracle import zipvoice_oracle as O
    from zipvoice_b200 import synth
    from zipvoice_b200.model import install
    from zipvoice_b200.solver import B200DistillEulerSolver, B200EulerSolver
    cfg = synth.small_model_cfg()
    ZipVoice = type("ZipVoice", (_RefShaped,), {})
    ZipVoiceDistill = type("ZipVoiceDistill", (_RefShaped,), {})
    ZipVoiceDialog = type("ZipVoiceDialog", (_RefShaped,), {})
    # ---- ZipVoice: sample() through the patched seam vs the reference golden and vs ZipVoiceB200
    sd = synth.make_state_dict(cfg, seed=1, stress=True)
    ref = ZipVoice(cfg, sd, synth.PAD_ID)
    eng = install(ref, cfg)
    assert type(ref.solver) is B200EulerSolver and not eng.distill and not eng.dialog
    tokens, prompt_tokens = synth.make_tokens(2, n_prompt=9, n_text=21, seed=5, ragged=True)
    pf, pfl = synth.make_prompt_features(2, 14, seed=6, ragged=True)
    ptl, tl = torch.tensor([len(t) for t in prompt_tokens]), torch.tensor([len(t) for t in tokens])
    T = int(O.predict_features_lens(pfl, ptl, tl, 1.3).max())
    torch.manual_seed(666)
    x0 = torch.randn(2, T, 100)
    out, wo_lens = ref.sample(tokens, prompt_tokens, pf, pfl, speed=1.3, t_shift=0.5, num_step=3, guidance_scale=1.0, x0=x0)
    g = _g("sample_small_stress.npz")
    assert torch.equal(wo_lens, g["wo_lens"])
    r = rel_l2(out, g["x1_wo"])
    assert r < TOL_MEL, f"patched reference-shaped sample() vs reference golden: {r:.3e}"
    m, _, _ = engine("zipvoice", "small", 1, True)
    mine = m.sample(tokens, prompt_tokens, pf, pfl, speed=1.3, t_shift=0.5, duration="predict", num_step=3,
                    guidance_scale=1.0, x0=x0.cuda())[0].cpu()
    assert rel_l2(out, mine) < TOL_MEL
    # ---- ZipVoiceDistill: the distill solver (guidance embedded) vs the reference golden
    sdd = synth.make_state_dict(cfg, seed=2, distill=True, stress=True)
    refd = ZipVoiceDistill(cfg, sdd, synth.PAD_ID)
    engd = install(refd, cfg)
    assert type(refd.solver) is B200DistillEulerSolver and engd.distill
    xs, text, speech, mask = case_inputs(2, 37, 11, seed=10)
    x1 = refd.solver.sample(x=xs, text_condition=text, speech_condition=speech, padding_mask=mask, num_step=2,
                            guidance_scale=3.0, t_shift=0.5).cpu()
    assert valid_rel_l2(x1, _g("euler_small_distill.npz")["x1"], mask) < TOL_MEL
    # ---- ZipVoiceDialog: speaker ids travel, the engine's text encoder adds the speaker rows
    sdg = synth.make_state_dict(cfg, seed=3, dialog=True, stress=True)
    refg = ZipVoiceDialog(cfg, sdg, synth.PAD_ID, dialog=(synth.SPK_A_ID, synth.SPK_B_ID))
    engg = install(refg, cfg)
    assert engg.dialog and type(refg.solver) is B200EulerSolver
    for e in (eng, engd, engg):
        e.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_second_engine_on_another_device_in_the_same_process():
    """function attributes (the > 48 KB dynamic shared-memory opt-in) are per DEVICE.  A handle created on
    cuda:1 after cuda:0 has run must launch the same kernels and return the same numbers."""
    from zipvoice_b200 import synth
    from zipvoice_b200.model import ZipVoiceB200
    m0, cfg, sd = engine("zipvoice", "small", 1, True)
    x0, text, speech, mask = case_inputs(2, 37, 11, seed=10)
    v0 = m0.forward_fm_decoder(torch.tensor(0.3), x0, text, speech, mask).cpu()
    m1 = ZipVoiceB200(cfg, sd, synth.VOCAB_SIZE, synth.PAD_ID, kind="zipvoice", device="cuda:1")
    v1 = m1.forward_fm_decoder(torch.tensor(0.3), x0, text, speech, mask)
    assert v1.device == torch.device("cuda:1")
    assert torch.equal(v1.cpu(), v0)
    x1 = m1.solver.sample(x=x0, text_condition=text, speech_condition=speech, padding_mask=mask, num_step=4,
                          guidance_scale=1.0, t_shift=0.5).cpu()
    assert valid_rel_l2(x1, _g("euler_small_stress.npz")["x1_cfg"], mask) < TOL_MEL
    assert torch.equal(m0.forward_fm_decoder(torch.tensor(0.3), x0, text, speech, mask).cpu(), v0)
    m1.engine.close()

"""Seeded synthetic weights and inputs for the ZipVoice hot path.

There is no network for checkpoints or datasets, so benchmarks, tests and golden
vectors all use tensors generated here.  ``make_state_dict`` produces a state dict
with exactly the reference's tensor names and shapes (the golden
script loads it into the unmodified reference with ``strict=True``), drawn with
init-like scales (reference zipformer.py:1108-1113,1141-1143,1345-1350,1416-1424,
1481-1483,1630-1636; scaling.py:460-480).  Every tensor is drawn from its own
generator seeded by (seed, crc32(name)) so values do not depend on creation order.
"""
import math
import zlib
from typing import Dict, List

import torch

# egs/zipvoice/conf/zipvoice_base.json ("model" section) -- the 123 M configuration.
BASE_MODEL_CFG = {
    "fm_decoder_downsampling_factor": [1, 2, 4, 2, 1],
    "fm_decoder_num_layers": [2, 2, 4, 4, 4],
    "fm_decoder_cnn_module_kernel": [31, 15, 7, 15, 31],
    "fm_decoder_feedforward_dim": 1536,
    "fm_decoder_num_heads": 4,
    "fm_decoder_dim": 512,
    "text_encoder_num_layers": 4,
    "text_encoder_feedforward_dim": 512,
    "text_encoder_cnn_module_kernel": 9,
    "text_encoder_num_heads": 4,
    "text_encoder_dim": 192,
    "query_head_dim": 32,
    "value_head_dim": 12,
    "pos_head_dim": 4,
    "pos_dim": 48,
    "time_embed_dim": 192,
    "text_embed_dim": 192,
    "feat_dim": 100,
}
VOCAB_SIZE = 364  # >= dialog speaker ids 360/361 (zipvoice_dialog.py:53-54)
PAD_ID = 0
SPK_A_ID, SPK_B_ID = 360, 361


def small_model_cfg(**over) -> dict:
    """A reduced-depth copy of the base config (same widths, so the same kernels run)."""
    cfg = dict(BASE_MODEL_CFG)
    cfg.update(fm_decoder_downsampling_factor=[1, 2, 1], fm_decoder_num_layers=[1, 1, 1],
               fm_decoder_cnn_module_kernel=[31, 15, 31], text_encoder_num_layers=1)
    cfg.update(over)
    return cfg


def _gen(seed: int, name: str) -> torch.Generator:
    return torch.Generator().manual_seed((seed * 1000003 + zlib.crc32(name.encode())) & 0x7FFFFFFF)


def _uniform(g, shape, a):
    return (torch.rand(shape, generator=g) * 2 - 1) * a


def _zipformer_shapes(p: str, in_dim, out_dim, D, F, H, dq, dp, dv, pos_dim, ds_list, nl_list, k_list,
                      time_dim, guidance: bool) -> List[tuple]:
    """(name, shape, kind, scale) for one TTSZipformer (zipformer.py:109-240)."""
    out = [(f"{p}.in_proj", (D, in_dim), "linear", 1.0), (f"{p}.out_proj", (out_dim, D), "linear", 1.0)]
    if time_dim > 0:
        out += [(f"{p}.time_embed.0", (2 * time_dim, time_dim), "linear", 1.0),
                (f"{p}.time_embed.2", (time_dim, 2 * time_dim), "linear", 1.0)]
        if guidance:
            out.append((f"{p}.guidance_scale_embed", (time_dim, time_dim), "scaled_nobias", 0.1))
    for i, (ds, nl, k) in enumerate(zip(ds_list, nl_list, k_list)):
        e = f"{p}.encoders.{i}" + (".encoder" if ds != 1 else "")
        if ds != 1:
            out.append((f"{p}.encoders.{i}.downsample.bias", (ds,), "ds_bias", 0.0))
            out.append((f"{p}.encoders.{i}.out_combiner.bypass_scale", (D,), "bypass", 0.5))
        if time_dim > 0:
            out.append((f"{e}.time_emb.1", (D, time_dim), "linear", 1.0))
        for j in range(nl):
            l = f"{e}.layers.{j}"
            out += [
                (f"{l}.bypass.bypass_scale", (D,), "bypass", 0.5),
                (f"{l}.bypass_mid.bypass_scale", (D,), "bypass", 0.5),
                (f"{l}.self_attn_weights.in_proj", (H * (2 * dq + dp), D), "scaled", dq ** -0.25),
                (f"{l}.self_attn_weights.linear_pos", (H * dp, pos_dim), "scaled_nobias", 0.05),
                (f"{l}.nonlin_attention.in_proj", (3 * (3 * D // 4), D), "linear", 1.0),
                (f"{l}.nonlin_attention.out_proj", (D, 3 * D // 4), "scaled", 0.05),
                (f"{l}.norm.log_scale", (), "log_scale", 1.0),
                (f"{l}.norm.bias", (D,), "norm_bias", 1e-4),
            ]
            for n in (1, 2):
                out += [(f"{l}.self_attn{n}.in_proj", (H * dv, D), "linear", 1.0),
                        (f"{l}.self_attn{n}.out_proj", (D, H * dv), "scaled", 0.05),
                        (f"{l}.conv_module{n}.in_proj", (2 * D, D), "linear", 1.0),
                        (f"{l}.conv_module{n}.depthwise_conv", (D, 1, k), "dwconv", 1.0),
                        (f"{l}.conv_module{n}.out_proj", (D, D), "scaled", 0.05)]
            for n, f in ((1, F * 3 // 4), (2, F), (3, F * 5 // 4)):
                out += [(f"{l}.feed_forward{n}.in_proj", (f, D), "linear", 1.0),
                        (f"{l}.feed_forward{n}.out_proj", (D, f), "scaled", 0.1)]
    return out


def make_state_dict(model_cfg: dict = None, seed: int = 0, distill: bool = False, dialog: bool = False,
                    stress: bool = False, vocab_size: int = VOCAB_SIZE, stereo: bool = False) -> Dict[str, torch.Tensor]:
    """fp32 CPU state dict with the reference's names.  ``stress`` scales up the attention and
    position projections, out-projections, bypass scales and BiasNorm biases so that every branch
    of the layer contributes visibly to the output."""
    c = model_cfg or BASE_MODEL_CFG
    common = dict(H=None, dq=c["query_head_dim"], dp=c["pos_head_dim"], dv=c["value_head_dim"], pos_dim=c["pos_dim"])
    specs = _zipformer_shapes(
        "fm_decoder", 3 * c["feat_dim"], c["feat_dim"], c["fm_decoder_dim"], c["fm_decoder_feedforward_dim"],
        c["fm_decoder_num_heads"], common["dq"], common["dp"], common["dv"], common["pos_dim"],
        c["fm_decoder_downsampling_factor"], c["fm_decoder_num_layers"], c["fm_decoder_cnn_module_kernel"],
        c["time_embed_dim"], distill)
    specs += _zipformer_shapes(
        "text_encoder", c["text_embed_dim"], c["feat_dim"], c["text_encoder_dim"], c["text_encoder_feedforward_dim"],
        c["text_encoder_num_heads"], common["dq"], common["dp"], common["dv"], common["pos_dim"],
        [1], [c["text_encoder_num_layers"]], [c["text_encoder_cnn_module_kernel"]], -1, False)
    sd: Dict[str, torch.Tensor] = {}
    for name, shape, kind, scale in specs:
        g = _gen(seed, name)
        if kind in ("linear", "scaled", "scaled_nobias"):
            fan_in = shape[1]
            s = scale
            if stress:
                if name.endswith("self_attn_weights.in_proj"):
                    s *= 4.0
                elif name.endswith("linear_pos"):
                    s *= 20.0
                elif name.endswith("out_proj") and "layers" in name:
                    s *= 4.0
            sd[name + ".weight"] = _uniform(g, shape, s / math.sqrt(fan_in))
            if kind == "linear":
                sd[name + ".bias"] = _uniform(g, (shape[0],), 1.0 / math.sqrt(fan_in))
            elif kind == "scaled":
                sd[name + ".bias"] = _uniform(g, (shape[0],), 0.1 * s)
        elif kind == "dwconv":
            a = 1.0 / math.sqrt(shape[2])
            sd[name + ".weight"] = _uniform(g, shape, a)
            sd[name + ".bias"] = _uniform(g, (shape[0],), a)
        elif kind == "bypass":
            sd[name] = torch.rand(shape, generator=g) * 0.8 + 0.2 if stress else torch.full(shape, 0.5)
        elif kind == "log_scale":
            sd[name] = (torch.rand((), generator=g) - 0.5) + 1.0 if stress else torch.tensor(1.0)
        elif kind == "norm_bias":
            sd[name] = torch.randn(shape, generator=g) * (0.5 if stress else 1e-4)
        elif kind == "ds_bias":
            sd[name] = torch.randn(shape, generator=g) if stress else torch.zeros(shape)
        else:
            raise AssertionError(kind)
    if stereo:
        # ZipVoiceDialogStereo: TTSZipformerTwoStream keeps two projection pairs in ModuleLists
        # (zipformer_two_stream.py:139-150): index 0 = two channels (5F -> D, D -> 2F), index 1 = mono
        F_, D_ = c["feat_dim"], c["fm_decoder_dim"]
        for k in ("weight", "bias"):
            sd[f"fm_decoder.in_proj.1.{k}"] = sd.pop(f"fm_decoder.in_proj.{k}")
            sd[f"fm_decoder.out_proj.1.{k}"] = sd.pop(f"fm_decoder.out_proj.{k}")
        for name, shape in (("fm_decoder.in_proj.0", (D_, 5 * F_)), ("fm_decoder.out_proj.0", (2 * F_, D_))):
            g = _gen(seed, name)
            sd[name + ".weight"] = _uniform(g, shape, 1.0 / math.sqrt(shape[1]))
            sd[name + ".bias"] = _uniform(g, (shape[0],), 1.0 / math.sqrt(shape[1]))
    sd["embed.weight"] = torch.randn(vocab_size, c["text_embed_dim"], generator=_gen(seed, "embed.weight"))
    if dialog:
        sd["spk_embed.weight"] = torch.randn(2, c["feat_dim"], generator=_gen(seed, "spk_embed.weight")) * 0.1
    return sd


# ---------------------------------------------------------------- synthetic inputs
def make_tokens(batch: int, n_prompt: int = 45, n_text: int = 150, seed: int = 1234, dialog: bool = False,
                ragged: bool = False):
    """Uniform ids in [1, 360); dialog inserts [S1]/[S2] alternately every ~25 tokens, text starts with [S1]."""
    g = torch.Generator().manual_seed(seed)
    prompts, texts = [], []
    for b in range(batch):
        np_, nt = n_prompt, n_text
        if ragged:
            np_ = max(4, int(n_prompt * (0.6 + 0.4 * torch.rand((), generator=g).item())))
            nt = max(4, int(n_text * (0.5 + 0.5 * torch.rand((), generator=g).item())))
        p = torch.randint(1, 360, (np_,), generator=g).tolist()
        t = torch.randint(1, 360, (nt,), generator=g).tolist()
        if dialog:
            for seq in (p, t):
                for k, pos in enumerate(range(0, len(seq), 25)):
                    seq[pos] = SPK_A_ID if k % 2 == 0 else SPK_B_ID
        prompts.append(p)
        texts.append(t)
    return texts, prompts


def make_prompt_features(batch: int, n_frames: int, feat_dim: int = 100, seed: int = 4321, ragged: bool = False):
    """Log-mel-like prompt features already scaled by feat_scale=0.1 (infer_zipvoice.py:349)."""
    g = torch.Generator().manual_seed(seed)
    feats = (torch.randn(batch, n_frames, feat_dim, generator=g) * 0.5 - 0.5)
    lens = torch.full((batch,), n_frames, dtype=torch.int64)
    if ragged:
        lens = (n_frames * (0.6 + 0.4 * torch.rand(batch, generator=g))).to(torch.int64).clamp(min=2)
        lens[0] = n_frames
        feats = feats * (torch.arange(n_frames)[None, :, None] < lens[:, None, None])
    return feats, lens


def make_noise(batch: int, n_frames: int, feat_dim: int = 100, seed: int = 666):
    return torch.randn(batch, n_frames, feat_dim, generator=torch.Generator().manual_seed(seed))


def make_audio(batch: int, n_samples: int, seed: int = 7):
    """0.1*randn low-passed with a 5-tap box filter plus two sines; 24 kHz mono."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(batch, n_samples + 4, generator=g) * 0.1
    x = x.unfold(1, 5, 1).mean(-1)
    t = torch.arange(n_samples) / 24000.0
    return x + 0.05 * torch.sin(2 * math.pi * 220.0 * t) + 0.03 * torch.sin(2 * math.pi * 1760.0 * t)


def make_vocos_state_dict(seed: int = 0) -> Dict[str, torch.Tensor]:
    """Seeded random weights of the Vocos mel-24khz architecture with the checkpoint's tensor names
    (``backbone.*``, ``head.out.*``): there is no network for the pretrained vocoder.  Scales are
    init-like but larger than the original trunc-normal(0.02) so that every block contributes;
    layer scale gamma ~ 1 / num_layers."""
    DIM, INTER, LAYERS, N_MELS, N_FFT = 512, 1536, 8, 100, 1024
    g = torch.Generator().manual_seed(1000 + seed)
    rn = lambda *s, std=1.0: torch.randn(*s, generator=g) * std
    sd = {"backbone.embed.weight": rn(DIM, N_MELS, 7, std=0.06), "backbone.embed.bias": rn(DIM, std=0.05),
          "backbone.norm.weight": 1 + rn(DIM, std=0.1), "backbone.norm.bias": rn(DIM, std=0.1),
          "backbone.final_layer_norm.weight": 1 + rn(DIM, std=0.1), "backbone.final_layer_norm.bias": rn(DIM, std=0.1),
          "head.out.weight": rn(N_FFT + 2, DIM, std=0.03), "head.out.bias": rn(N_FFT + 2, std=0.3)}
    for i in range(LAYERS):
        p = f"backbone.convnext.{i}"
        sd.update({f"{p}.dwconv.weight": rn(DIM, 1, 7, std=0.3), f"{p}.dwconv.bias": rn(DIM, std=0.1),
                   f"{p}.norm.weight": 1 + rn(DIM, std=0.1), f"{p}.norm.bias": rn(DIM, std=0.1),
                   f"{p}.pwconv1.weight": rn(INTER, DIM, std=0.05), f"{p}.pwconv1.bias": rn(INTER, std=0.1),
                   f"{p}.pwconv2.weight": rn(DIM, INTER, std=0.05), f"{p}.pwconv2.bias": rn(DIM, std=0.1),
                   f"{p}.gamma": torch.full((DIM,), 1.0 / LAYERS) * (1 + rn(DIM, std=0.2))})
    return sd


def make_intermediate_inputs(seed: int = 31, B: int = 2, T: int = 43, n_tokens: int = 17):
    """Seeded inputs of ZipVoice.sample_intermediate (zipvoice.py:488-534): ragged feature lengths, a prompt-like
    speech-condition mask (True = not conditioned) and a per-sample guidance tensor (batch, 1, 1)."""
    g = torch.Generator().manual_seed(seed)
    tokens = [torch.randint(1, VOCAB_SIZE, (n_tokens - 3 * b,), generator=g).tolist() for b in range(B)]
    features = torch.randn(B, T, 100, generator=g) * 0.5 - 0.5
    features_lens = torch.tensor([T - 6 * b for b in range(B)], dtype=torch.int64)
    noise = torch.randn(B, T, 100, generator=g)
    start = torch.tensor([9 + 4 * b for b in range(B)])
    speech_condition_mask = torch.arange(T)[None, :] >= start[:, None]
    guidance_scale = torch.tensor([0.5 + 1.5 * b for b in range(B)]).reshape(B, 1, 1)
    return {"tokens": tokens, "features": features, "features_lens": features_lens, "noise": noise,
            "speech_condition_mask": speech_condition_mask, "guidance_scale": guidance_scale}

"""TEST INFRASTRUCTURE ONLY -- torch-fp32 CPU restatement of the ZipVoice sampling path:
EulerSolver / DiffusionModel (CFG) / DistillDiffusionModel, ZipVoice.forward_fm_decoder,
the text-condition helpers and ZipVoice.sample, plus the dialog speaker embedding.
Citations are relative to /root/reference/zipvoice/.  Parity pinned by tests/golden.
"""
import math
from typing import Dict, List, Optional, Union

import torch

from . import zipformer_oracle as zo

Tensor = torch.Tensor


def fm_cfg(model_cfg: dict) -> dict:
    return dict(downsampling_factor=model_cfg["fm_decoder_downsampling_factor"],
                num_layers=model_cfg["fm_decoder_num_layers"], num_heads=model_cfg["fm_decoder_num_heads"],
                query_head_dim=model_cfg["query_head_dim"], pos_head_dim=model_cfg["pos_head_dim"],
                pos_dim=model_cfg["pos_dim"], time_embed_dim=model_cfg["time_embed_dim"])


def text_cfg(model_cfg: dict) -> dict:
    return dict(downsampling_factor=[1], num_layers=[model_cfg["text_encoder_num_layers"]],
                num_heads=model_cfg["text_encoder_num_heads"], query_head_dim=model_cfg["query_head_dim"],
                pos_head_dim=model_cfg["pos_head_dim"], pos_dim=model_cfg["pos_dim"], time_embed_dim=-1)


def get_time_steps(t_start=0.0, t_end=1.0, num_step=10, t_shift=1.0) -> Tensor:
    """models/modules/solver.py:256-281."""
    ts = torch.linspace(t_start, t_end, num_step + 1)
    return t_shift * ts / (1 + (t_shift - 1) * ts)


def forward_fm_decoder(sd, model_cfg, t: Tensor, xt, text_condition, speech_condition, padding_mask=None,
                       guidance_scale: Optional[Tensor] = None, taps=None) -> Tensor:
    """models/zipvoice.py:135-185."""
    x = torch.cat([xt, text_condition, speech_condition], dim=2)
    assert t.dim() in (0, 3)
    while t.dim() > 1 and t.size(-1) == 1:
        t = t.squeeze(-1)
    if t.dim() == 0:
        t = t.repeat(x.shape[0])
    if guidance_scale is not None:
        while guidance_scale.dim() > 1 and guidance_scale.size(-1) == 1:
            guidance_scale = guidance_scale.squeeze(-1)
        if guidance_scale.dim() == 0:
            guidance_scale = guidance_scale.repeat(x.shape[0])
    return zo.tts_zipformer(sd, "fm_decoder", fm_cfg(model_cfg), x, t, padding_mask, guidance_scale, taps)


def cfg_velocity(sd, model_cfg, t: Tensor, x, text_condition, speech_condition, padding_mask,
                 guidance_scale: Union[float, Tensor] = 0.0) -> Tensor:
    """models/modules/solver.py:40-110 DiffusionModel.forward (CFG batch doubling)."""
    if not torch.is_tensor(guidance_scale):
        guidance_scale = torch.tensor(guidance_scale, dtype=t.dtype)
    if (guidance_scale == 0.0).all():
        return forward_fm_decoder(sd, model_cfg, t, x, text_condition, speech_condition, padding_mask)
    assert t.dim() == 0
    x2 = torch.cat([x] * 2, dim=0)
    m2 = torch.cat([padding_mask] * 2, dim=0)
    text2 = torch.cat([torch.zeros_like(text_condition), text_condition], dim=0)
    if t > 0.5:
        speech2 = torch.cat([torch.zeros_like(speech_condition), speech_condition], dim=0)
    else:
        guidance_scale = guidance_scale * 2
        speech2 = torch.cat([speech_condition, speech_condition], dim=0)
    u, c = forward_fm_decoder(sd, model_cfg, t, x2, text2, speech2, m2).chunk(2, dim=0)
    return (1 + guidance_scale) * c - guidance_scale * u


def distill_velocity(sd, model_cfg, t, x, text_condition, speech_condition, padding_mask, guidance_scale=0.0):
    """models/modules/solver.py:127-165 DistillDiffusionModel.forward (guidance is an embedding input)."""
    if not torch.is_tensor(guidance_scale):
        guidance_scale = torch.tensor(guidance_scale, dtype=t.dtype)
    return forward_fm_decoder(sd, model_cfg, t, x, text_condition, speech_condition, padding_mask, guidance_scale)


def euler_sample(sd, model_cfg, x, text_condition, speech_condition, padding_mask, num_step=10,
                 guidance_scale=0.0, t_start=0.0, t_end=1.0, t_shift=1.0, distill=False,
                 velocities: Optional[list] = None) -> Tensor:
    """models/modules/solver.py:182-240 EulerSolver.sample."""
    assert isinstance(t_start, float) and isinstance(t_end, float)
    ts = get_time_steps(t_start, t_end, num_step, t_shift)
    fn = distill_velocity if distill else cfg_velocity
    for step in range(num_step):
        v = fn(sd, model_cfg, ts[step], x, text_condition, speech_condition, padding_mask, guidance_scale)
        if velocities is not None:
            velocities.append(v)
        x = x + v * (ts[step + 1] - ts[step])
    return x


# ---- text-condition helpers (utils/common.py) ----
def pad_labels(y: List[List[int]], pad_id: int) -> Tensor:
    """utils/common.py:255-268 (one pad token is APPENDED to every row, then right-pad)."""
    y = [list(t) + [pad_id] for t in y]
    n = max(len(t) for t in y)
    return torch.tensor([t + [pad_id] * (n - len(t)) for t in y], dtype=torch.int64)


def make_pad_mask(lengths: Tensor, max_len: int = 0) -> Tensor:
    """utils/common.py:395-420; True = padded."""
    max_len = max(max_len, int(lengths.max()))
    return torch.arange(max_len)[None, :] >= lengths[:, None]


def get_tokens_index(features_lens: Tensor, tokens_lens: Tensor, num_frames: int) -> Tensor:
    """utils/common.py:246-252 + 271-295: every token gets floor(T_i/ntok_i) frames, the remaining
    frames (and batch padding) index position ntok_i (the appended pad token)."""
    B = features_lens.numel()
    ans = torch.zeros(B, num_frames, dtype=torch.int64)
    for b in range(B):
        n = int(tokens_lens[b])
        d = int(features_lens[b]) // n
        cur = 0
        for i in range(n):
            ans[b, cur:cur + d] = i
            cur += d
        ans[b, cur:] = n
    return ans


def extract_spk_masks(tokens_padded: Tensor, pad_id: int, spk_a_id: int, spk_b_id: int):
    """models/zipvoice_dialog.py:118-125: parity of the running count of [S1]/[S2] tokens; pads -> -1."""
    turn = ((tokens_padded == spk_a_id) | (tokens_padded == spk_b_id)).long().cumsum(dim=1) % 2
    return torch.where(tokens_padded == pad_id, -1, turn)


def forward_text_embed(sd, model_cfg, tokens: List[List[int]], pad_id=0, dialog=None):
    """models/zipvoice.py:187-212; dialog=(spk_a_id, spk_b_id) adds models/zipvoice_dialog.py:127-159."""
    tp = pad_labels(tokens, pad_id)
    embed = sd["embed.weight"][tp]
    tokens_lens = torch.tensor([len(t) for t in tokens], dtype=torch.int64)
    mask = make_pad_mask(tokens_lens, embed.shape[1])
    embed = zo.tts_zipformer(sd, "text_encoder", text_cfg(model_cfg), embed, None, mask)
    if dialog is not None:
        spk = extract_spk_masks(tp, pad_id, dialog[0], dialog[1])
        embed = embed.clone()
        embed[spk == 0] += sd["spk_embed.weight"][0]
        embed[spk == 1] += sd["spk_embed.weight"][1]
    return embed, tokens_lens


def forward_text_condition(embed: Tensor, tokens_lens: Tensor, features_lens: Tensor):
    """models/zipvoice.py:214-251."""
    T = int(features_lens.max())
    padding_mask = make_pad_mask(features_lens, T)
    idx = get_tokens_index(features_lens, tokens_lens, T)
    text_condition = torch.gather(embed, 1, idx.unsqueeze(-1).expand(embed.size(0), T, embed.size(-1)))
    return text_condition, padding_mask


def predict_features_lens(prompt_features_lens, prompt_tokens_lens, tokens_lens, speed: float) -> Tensor:
    """models/zipvoice.py:323-325."""
    return prompt_features_lens + torch.ceil(
        prompt_features_lens / prompt_tokens_lens * tokens_lens / speed).to(torch.int64)


def sample(sd, model_cfg, tokens, prompt_tokens, prompt_features, prompt_features_lens, features_lens=None,
           speed=1.0, t_shift=1.0, duration="predict", num_step=5, guidance_scale=0.5, distill=False,
           pad_id=0, dialog=None, x0: Optional[Tensor] = None):
    """models/zipvoice.py:388-486 ZipVoice.sample.  x0 may be injected (the reference draws it from
    the global torch RNG, :453-458)."""
    assert duration in ["real", "predict"]
    cat_tokens = [p + t for p, t in zip(prompt_tokens, tokens)]
    embed, cat_lens = forward_text_embed(sd, model_cfg, cat_tokens, pad_id, dialog)
    if duration == "predict":
        ptl = torch.tensor([len(t) for t in prompt_tokens], dtype=torch.int64)
        tl = torch.tensor([len(t) for t in tokens], dtype=torch.int64)
        total_lens = predict_features_lens(prompt_features_lens, ptl, tl, speed)
    else:
        assert features_lens is not None
        total_lens = prompt_features_lens + features_lens
    text_condition, padding_mask = forward_text_condition(embed, cat_lens, total_lens)
    B, T, _ = text_condition.shape
    speech_condition = torch.nn.functional.pad(prompt_features, (0, 0, 0, T - prompt_features.size(1)))
    sc_mask = make_pad_mask(prompt_features_lens, T)
    speech_condition = torch.where(sc_mask.unsqueeze(-1), torch.zeros_like(speech_condition), speech_condition)
    if x0 is None:
        x0 = torch.randn(B, T, prompt_features.size(-1))
    x1 = euler_sample(sd, model_cfg, x0, text_condition, speech_condition, padding_mask, num_step=num_step,
                      guidance_scale=guidance_scale, t_shift=t_shift, distill=distill)
    wo_lens = (~padding_mask).sum(-1) - prompt_features_lens
    x1_prompt = torch.zeros(B, int(prompt_features_lens.max()), x1.size(2))
    x1_wo = torch.zeros(B, int(wo_lens.max()), x1.size(2))
    for i in range(B):
        pl, wl = int(prompt_features_lens[i]), int(wo_lens[i])
        x1_wo[i, :wl] = x1[i, pl:pl + wl]
        x1_prompt[i, :pl] = x1[i, :pl]
    return x1_wo, wo_lens, x1_prompt, prompt_features_lens


@torch.inference_mode()
def sample_intermediate(sd, model_cfg, tokens, features, features_lens, noise, speech_condition_mask, t_start: float,
                        t_end: float, num_step: int = 1, guidance_scale=None, distill=False, pad_id=0):
    """models/zipvoice.py:488-534 ZipVoice.sample_intermediate (teacher roll-out of the distillation recipe,
    bin/train_zipvoice_distill.py:439-460): real feature lengths, speech condition = features outside the mask,
    Euler steps over [t_start, t_end] without time shifting, guidance (batch, 1, 1) per sample."""
    embed, tokens_lens = forward_text_embed(sd, model_cfg, tokens, pad_id)
    text_condition, padding_mask = forward_text_condition(embed, tokens_lens, features_lens)
    speech_condition = torch.where(speech_condition_mask.unsqueeze(-1), torch.zeros(()), features)
    x = euler_sample(sd, model_cfg, noise, text_condition, speech_condition, padding_mask, num_step=num_step,
                     guidance_scale=guidance_scale, t_start=t_start, t_end=t_end, distill=distill)
    return x, (~padding_mask).sum(-1)

"""Host-side mirror of the reference model API (zipvoice/models/zipvoice.py) on top of the engine.

``ZipVoiceB200`` offers ``sample`` / ``sample_intermediate`` / ``forward_fm_decoder`` /
``forward_text_embed`` / ``forward_text_condition`` with the reference's argument names and return
values, so ``infer_zipvoice.generate_sentence(model=...)`` can take it unchanged.  ``install`` patches
an existing reference ``ZipVoice`` module in place instead.
"""
from typing import List, Optional

import torch

from .engine import Engine
from .solver import B200DistillEulerSolver, B200EulerSolver


def pad_labels(y: List[List[int]], pad_id: int, device) -> torch.Tensor:
    """utils/common.py:255-268 (one pad token appended to every row)."""
    y = [list(t) + [pad_id] for t in y]
    n = max(len(t) for t in y)
    return torch.tensor([t + [pad_id] * (n - len(t)) for t in y], dtype=torch.int64, device=device)


class ZipVoiceB200:
    """kind: "zipvoice" | "distill" | "dialog" | "dialog_stereo"
    (ZipVoice / ZipVoiceDistill / ZipVoiceDialog / ZipVoiceDialogStereo, zipvoice_dialog.py:218-256: two-channel
    features of width 2 * feat_dim through the first stream of TTSZipformerTwoStream)."""

    def __init__(self, model_cfg: dict, state_dict, vocab_size: int, pad_id: int = 0, kind: str = "zipvoice",
                 spk_a_id: int = 360, spk_b_id: int = 361, device="cuda:0"):
        assert kind in ("zipvoice", "distill", "dialog", "dialog_stereo")
        self.engine = Engine(model_cfg, state_dict, vocab_size, pad_id, distill=(kind == "distill"),
                             dialog=kind.startswith("dialog"), spk_a_id=spk_a_id, spk_b_id=spk_b_id, device=device,
                             stereo=(kind == "dialog_stereo"))
        self.kind, self.pad_id, self.feat_dim = kind, pad_id, model_cfg["feat_dim"]
        self.device = self.engine.device
        self.solver = (B200DistillEulerSolver if kind == "distill" else B200EulerSolver)(self.engine)

    def eval(self):
        return self

    def to(self, device):
        assert torch.device(device) == self.device, "a handle is bound to the device it was created on"
        return self

    # ---- zipvoice.py:135-185
    def forward_fm_decoder(self, t, xt, text_condition, speech_condition, padding_mask=None, guidance_scale=None):
        t = torch.as_tensor(t)
        assert t.dim() in (0, 3)  # zipvoice.py:165: a scalar, or (N, 1, 1) with one time step per row
        return self.engine.forward_fm_decoder(t, xt, text_condition, speech_condition, padding_mask, guidance_scale)

    # ---- zipvoice.py:187-212 (+ zipvoice_dialog.py:127-159 inside the library)
    def forward_text_embed(self, tokens: List[List[int]]):
        tp = pad_labels(tokens, self.pad_id, self.device)
        tokens_lens = torch.tensor([len(t) for t in tokens], dtype=torch.int64, device=self.device)
        return self.engine.text_embed(tp, tokens_lens), tokens_lens

    # ---- zipvoice.py:214-251
    def forward_text_condition(self, embed, tokens_lens, features_lens):
        num_frames = int(features_lens.max())
        return self.engine.text_condition(embed, tokens_lens, features_lens, num_frames)

    def forward_text_train(self, tokens, features_lens):
        embed, tokens_lens = self.forward_text_embed(tokens)
        return self.forward_text_condition(embed, tokens_lens, features_lens)

    def forward_text_inference_gt_duration(self, tokens, features_lens, prompt_tokens, prompt_features_lens):
        tokens = [p + t for p, t in zip(prompt_tokens, tokens)]
        embed, tokens_lens = self.forward_text_embed(tokens)
        return self.forward_text_condition(embed, tokens_lens, prompt_features_lens + features_lens)

    def forward_text_inference_ratio_duration(self, tokens, prompt_tokens, prompt_features_lens, speed):
        cat = [p + t for p, t in zip(prompt_tokens, tokens)]
        ptl = torch.tensor([len(t) for t in prompt_tokens], dtype=torch.int64, device=self.device)
        tl = torch.tensor([len(t) for t in tokens], dtype=torch.int64, device=self.device)
        embed, cat_lens = self.forward_text_embed(cat)
        pfl = prompt_features_lens.to(self.device)
        features_lens = pfl + torch.ceil(pfl / ptl * tl / speed).to(dtype=torch.int64)  # zipvoice.py:323-325
        return self.forward_text_condition(embed, cat_lens, features_lens)

    # ---- zipvoice.py:388-486
    @torch.inference_mode()
    def sample(self, tokens, prompt_tokens, prompt_features, prompt_features_lens, features_lens=None, speed=1.0,
               t_shift=1.0, duration="predict", num_step=5, guidance_scale=0.5, x0: Optional[torch.Tensor] = None):
        assert duration in ["real", "predict"]
        prompt_features = prompt_features.to(self.device, torch.float32)
        prompt_features_lens = prompt_features_lens.to(self.device)
        if duration == "predict":
            text_condition, padding_mask = self.forward_text_inference_ratio_duration(
                tokens, prompt_tokens, prompt_features_lens, speed)
        else:
            assert features_lens is not None
            text_condition, padding_mask = self.forward_text_inference_gt_duration(
                tokens, features_lens.to(self.device), prompt_tokens, prompt_features_lens)
        B, T, _ = text_condition.shape
        speech_condition = torch.nn.functional.pad(prompt_features, (0, 0, 0, T - prompt_features.size(1)))
        keep = torch.arange(T, device=self.device)[None, :] < prompt_features_lens[:, None]
        speech_condition = speech_condition * keep.unsqueeze(-1)
        if x0 is None:
            x0 = torch.randn(B, T, prompt_features.size(-1), device=self.device)  # zipvoice.py:453-458
        x1 = self.solver.sample(x=x0, text_condition=text_condition, speech_condition=speech_condition,
                                padding_mask=padding_mask, num_step=num_step, guidance_scale=guidance_scale,
                                t_shift=t_shift)
        # prompt strip without the per-utterance host loop of zipvoice.py:469-486: one gather
        wo_lens = (~padding_mask).sum(-1) - prompt_features_lens
        max_wo, max_p = int(wo_lens.max()), int(prompt_features_lens.max())
        ar = torch.arange(max(max_wo, 1), device=self.device)[None, :]
        idx = (prompt_features_lens[:, None] + ar).clamp(max=T - 1)
        x1_wo = torch.gather(x1, 1, idx.unsqueeze(-1).expand(B, idx.size(1), x1.size(2)))
        x1_wo = x1_wo * (ar < wo_lens[:, None]).unsqueeze(-1)
        x1_prompt = x1[:, :max_p] * keep[:, :max_p].unsqueeze(-1)
        return x1_wo[:, :max_wo], wo_lens, x1_prompt, prompt_features_lens

    # ---- zipvoice.py:488-534
    @torch.inference_mode()
    def sample_intermediate(self, tokens, features, features_lens, noise, speech_condition_mask, t_start, t_end,
                            num_step=1, guidance_scale=None):
        text_condition, padding_mask = self.forward_text_train(tokens, features_lens.to(self.device))
        speech_condition = torch.where(speech_condition_mask.to(self.device).unsqueeze(-1), 0,
                                       features.to(self.device))
        x = self.solver.sample(x=noise, text_condition=text_condition, speech_condition=speech_condition,
                               padding_mask=padding_mask, num_step=num_step, guidance_scale=guidance_scale,
                               t_start=t_start, t_end=t_end)
        return x, (~padding_mask).sum(-1)


def install(ref_model, model_cfg: dict, device="cuda:0") -> Engine:
    """Patch a reference ``ZipVoice`` / ``ZipVoiceDistill`` / ``ZipVoiceDialog`` nn.Module in place:
    weights are read from its state_dict and ``model.solver`` is replaced (zipvoice.py:133)."""
    name = type(ref_model).__name__
    distill, dialog, stereo = "Distill" in name, "Dialog" in name, "Stereo" in name
    sd = ref_model.state_dict()
    eng = Engine(model_cfg, sd, vocab_size=sd["embed.weight"].shape[0], pad_id=ref_model.pad_id, distill=distill,
                 dialog=dialog, spk_a_id=getattr(ref_model, "spk_a_id", 360),
                 spk_b_id=getattr(ref_model, "spk_b_id", 361), device=device, stereo=stereo)
    ref_model.solver = (B200DistillEulerSolver if distill else B200EulerSolver)(eng)
    return eng

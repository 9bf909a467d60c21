"""TEST INFRASTRUCTURE ONLY -- torch-fp32 CPU restatement of the TTSZipformer network
in eval mode.  Functional style: every function takes the reference's ``state_dict``
(same tensor names and shapes) plus a key prefix.

Each function cites the reference lines it follows (paths relative to
/root/reference/zipvoice/models/modules/).  Parity pinned by tests/golden (see
oracle/__init__.py).
"""
import math
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
SD = Dict[str, Tensor]


def swoosh_l(x: Tensor) -> Tensor:
    """scaling.py:1174-1180 SwooshLForward: log(1+exp(x-4)) - 0.08x - 0.035 (inf guard)."""
    xo = x - 4.0
    ls = (1.0 + xo.exp()).log()
    ls = torch.where(ls == float("inf"), xo, ls)
    return ls - 0.08 * x - 0.035


def swoosh_r(x: Tensor) -> Tensor:
    """scaling.py:1185-1191 SwooshRForward (same as logaddexp form :1116 in eval)."""
    xo = x - 1.0
    ls = (1.0 + xo.exp()).log()
    ls = torch.where(ls == float("inf"), xo, ls)
    return ls - 0.08 * x - 0.313261687


def linear(sd: SD, p: str, x: Tensor) -> Tensor:
    return F.linear(x, sd[p + ".weight"], sd.get(p + ".bias"))


def bias_norm(sd: SD, p: str, x: Tensor) -> Tensor:
    """scaling.py:345-348 / 439-442: x * mean((x-b)^2)^-0.5 * exp(log_scale)."""
    scales = torch.mean((x - sd[p + ".bias"]) ** 2, dim=-1, keepdim=True) ** -0.5
    return x * (scales * sd[p + ".log_scale"].exp())


def bypass(sd: SD, p: str, orig: Tensor, x: Tensor) -> Tensor:
    """zipformer.py:798-804 (eval: raw bypass_scale, :775-776)."""
    return orig + (x - orig) * sd[p + ".bypass_scale"]


def timestep_embedding(t: Tensor, dim: int, max_period: float = 10000.0) -> Tensor:
    """zipformer.py:47-69: [cos(t f_k) | sin(t f_k)], f_k = exp(-ln(1e4) k / half)."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(half, dtype=torch.float32) / half)
    args = t[..., None].float() * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)


def rel_pos_encoding(L: int, pos_dim: int) -> Tensor:
    """zipformer.py:983-1056 CompactRelPositionalEncoding -> (2L-1, pos_dim), row r
    holds relative offset r-(L-1)."""
    x = torch.arange(-(L - 1), L).to(torch.float32).unsqueeze(1)
    freqs = 1 + torch.arange(pos_dim // 2)
    c = pos_dim ** 0.5
    xc = c * x.sign() * ((x.abs() + c).log() - math.log(c))
    length_scale = pos_dim / (2.0 * math.pi)
    xa = (xc / length_scale).atan()
    pe = torch.zeros(x.shape[0], pos_dim)
    pe[:, 0::2] = (xa * freqs).cos()
    pe[:, 1::2] = (xa * freqs).sin()
    pe[:, -1] = 1.0
    return pe


def attention_weights(sd: SD, p: str, x: Tensor, pos_emb: Tensor, key_padding_mask: Optional[Tensor],
                      H: int, dq: int, dp: int) -> Tensor:
    """zipformer.py:1149-1306.  x (L,B,D) -> (H,B,L,L).  No 1/sqrt(d) scaling; relative
    position scores re-indexed col = L-1-i+j (:1239-1248); key padding -> -1000 (:1286-1289)."""
    L, B, _ = x.shape
    proj = linear(sd, p + ".in_proj", x)
    qd = H * dq
    q = proj[..., :qd].reshape(L, B, H, dq).permute(2, 1, 0, 3)
    k = proj[..., qd:2 * qd].reshape(L, B, H, dq).permute(2, 1, 3, 0)
    pq = proj[..., 2 * qd:].reshape(L, B, H, dp).permute(2, 1, 0, 3)
    scores = torch.matmul(q, k)
    pe = F.linear(pos_emb, sd[p + ".linear_pos.weight"])  # (2L-1, H*dp)
    pe = pe.reshape(1, 2 * L - 1, H, dp).permute(2, 0, 3, 1)  # (H,1,dp,2L-1)
    ps = torch.matmul(pq, pe)  # (H,B,L,2L-1)
    idx = (L - 1) - torch.arange(L)[:, None] + torch.arange(L)[None, :]
    ps = torch.gather(ps, 3, idx[None, None].expand(H, B, L, L))
    scores = scores + ps
    if key_padding_mask is not None:
        scores = scores.masked_fill(key_padding_mask.unsqueeze(1), -1000.0)
    return scores.softmax(dim=-1)


def self_attention(sd: SD, p: str, x: Tensor, attn: Tensor) -> Tensor:
    """zipformer.py:1359-1396."""
    L, B, _ = x.shape
    H = attn.shape[0]
    v = linear(sd, p + ".in_proj", x).reshape(L, B, H, -1).permute(2, 1, 0, 3)
    o = torch.matmul(attn, v).permute(2, 1, 0, 3).reshape(L, B, -1)
    return linear(sd, p + ".out_proj", o)


def nonlin_attention(sd: SD, p: str, x: Tensor, attn0: Tensor) -> Tensor:
    """zipformer.py:1499-1544 with attn0 = attn[0:1] (one head of width hidden, :542)."""
    L, B, _ = x.shape
    s, xx, y = linear(sd, p + ".in_proj", x).chunk(3, dim=2)
    xx = xx * torch.tanh(s)
    xx = xx.reshape(L, B, 1, -1).permute(2, 1, 0, 3)
    xx = torch.matmul(attn0, xx).permute(2, 1, 0, 3).reshape(L, B, -1)
    return linear(sd, p + ".out_proj", xx * y)


def feedforward(sd: SD, p: str, x: Tensor) -> Tensor:
    """zipformer.py:1433-1439 + scaling.py:1322-1334 (SwooshL then linear)."""
    return linear(sd, p + ".out_proj", swoosh_l(linear(sd, p + ".in_proj", x)))


def conv_module(sd: SD, p: str, x: Tensor, key_padding_mask: Optional[Tensor]) -> Tensor:
    """zipformer.py:1638-1680: GLU, zero padded frames, depthwise conv (same pad), SwooshR, linear."""
    xx, s = linear(sd, p + ".in_proj", x).chunk(2, dim=2)
    xx = (xx * torch.sigmoid(s)).permute(1, 2, 0)  # (B,C,L)
    if key_padding_mask is not None:
        xx = xx.masked_fill(key_padding_mask.unsqueeze(1).expand_as(xx), 0.0)
    w = sd[p + ".depthwise_conv.weight"]
    xx = F.conv1d(xx, w, sd[p + ".depthwise_conv.bias"], padding=w.shape[-1] // 2, groups=w.shape[0])
    xx = xx.permute(2, 0, 1)
    return linear(sd, p + ".out_proj", swoosh_r(xx))


def encoder_layer(sd: SD, p: str, src: Tensor, pos_emb: Tensor, time_emb: Optional[Tensor],
                  key_padding_mask: Optional[Tensor], H: int, dq: int, dp: int,
                  taps: Optional[dict] = None) -> Tensor:
    """zipformer.py:489-642 in eval mode (the 13 steps of SURVEY.md 3.3)."""
    orig = src
    attn = attention_weights(sd, p + ".self_attn_weights", src, pos_emb, key_padding_mask, H, dq, dp)
    if time_emb is not None:
        src = src + time_emb
    src = src + feedforward(sd, p + ".feed_forward1", src)
    src = src + nonlin_attention(sd, p + ".nonlin_attention", src, attn[0:1])
    src = src + self_attention(sd, p + ".self_attn1", src, attn)
    if time_emb is not None:
        src = src + time_emb
    src = src + conv_module(sd, p + ".conv_module1", src, key_padding_mask)
    src = src + feedforward(sd, p + ".feed_forward2", src)
    src = bypass(sd, p + ".bypass_mid", orig, src)
    src = src + self_attention(sd, p + ".self_attn2", src, attn)
    if time_emb is not None:
        src = src + time_emb
    src = src + conv_module(sd, p + ".conv_module2", src, key_padding_mask)
    src = src + feedforward(sd, p + ".feed_forward3", src)
    src = bias_norm(sd, p + ".norm", src)
    src = bypass(sd, p + ".bypass", orig, src)
    if taps is not None:
        taps[p] = src
    return src


def encoder_stack(sd: SD, p: str, n_layers: int, src: Tensor, time_emb: Optional[Tensor],
                  key_padding_mask: Optional[Tensor], H: int, dq: int, dp: int, pos_dim: int,
                  taps: Optional[dict] = None) -> Tensor:
    """zipformer.py:702-744 Zipformer2Encoder.forward."""
    pos_emb = rel_pos_encoding(src.shape[0], pos_dim)
    te = None
    if time_emb is not None:
        te = linear(sd, p + ".time_emb.1", swoosh_r(time_emb))  # (B,D), :677-680,729
    for j in range(n_layers):
        src = encoder_layer(sd, f"{p}.layers.{j}", src, pos_emb, te, key_padding_mask, H, dq, dp, taps)
    return src


def simple_downsample(sd: SD, p: str, src: Tensor, ds: int) -> Tensor:
    """zipformer.py:887-913: right-pad with last frame, weighted sum with softmax(bias)."""
    L, B, C = src.shape
    dL = (L + ds - 1) // ds
    pad = dL * ds - L
    src = torch.cat((src, src[L - 1:].expand(pad, B, C)), dim=0)
    w = sd[p + ".bias"].softmax(dim=0)
    return (src.reshape(dL, ds, B, C) * w[:, None, None]).sum(dim=1)


def downsampled_stack(sd: SD, p: str, n_layers: int, ds: int, src: Tensor, time_emb, key_padding_mask,
                      H, dq, dp, pos_dim, taps=None) -> Tensor:
    """zipformer.py:850-870 DownsampledZipformer2Encoder.forward."""
    orig = src
    x = simple_downsample(sd, p + ".downsample", src, ds)
    m = key_padding_mask[..., ::ds] if key_padding_mask is not None else None
    x = encoder_stack(sd, p + ".encoder", n_layers, x, time_emb, m, H, dq, dp, pos_dim, taps)
    L2, B, C = x.shape
    x = x.unsqueeze(1).expand(L2, ds, B, C).reshape(L2 * ds, B, C)[: orig.shape[0]]  # :925-935,868
    return bypass(sd, p + ".out_combiner", orig, x)


def tts_zipformer(sd: SD, p: str, cfg: dict, x: Tensor, t: Optional[Tensor], padding_mask: Optional[Tensor],
                  guidance_scale: Optional[Tensor] = None, taps: Optional[dict] = None) -> Tensor:
    """zipformer.py:242-293 TTSZipformer.forward.  x (B,T,Cin) -> (B,T,Cout).
    cfg: downsampling_factor, num_layers, num_heads, query_head_dim, pos_head_dim, pos_dim, time_embed_dim."""
    # TTSZipformerTwoStream (zipformer_two_stream.py:240-247): two projection pairs, chosen by the input width
    pj = ""
    if p + ".in_proj.weight" not in sd:
        pj = ".0" if x.size(2) == sd[p + ".in_proj.0.weight"].shape[1] else ".1"
        assert x.size(2) == sd[p + ".in_proj" + pj + ".weight"].shape[1], x.shape
    x = linear(sd, p + ".in_proj" + pj, x.permute(1, 0, 2))
    time_emb = None
    if t is not None:
        assert t.dim() == 1, t.shape
        time_emb = timestep_embedding(t, cfg["time_embed_dim"])
        if guidance_scale is not None:
            assert guidance_scale.dim() == 1
            time_emb = time_emb + F.linear(
                timestep_embedding(guidance_scale, cfg["time_embed_dim"]), sd[p + ".guidance_scale_embed.weight"])
        time_emb = linear(sd, p + ".time_embed.2", swoosh_r(linear(sd, p + ".time_embed.0", time_emb)))
    H, dq, dp, pos_dim = cfg["num_heads"], cfg["query_head_dim"], cfg["pos_head_dim"], cfg["pos_dim"]
    for i, (ds, nl) in enumerate(zip(cfg["downsampling_factor"], cfg["num_layers"])):
        sp = f"{p}.encoders.{i}"
        if ds == 1:
            x = encoder_stack(sd, sp, nl, x, time_emb, padding_mask, H, dq, dp, pos_dim, taps)
        else:
            x = downsampled_stack(sd, sp, nl, ds, x, time_emb, padding_mask, H, dq, dp, pos_dim, taps)
        if taps is not None:
            taps[sp] = x
    return linear(sd, p + ".out_proj" + pj, x).permute(1, 0, 2)

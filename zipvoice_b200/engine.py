"""Host-side owner of one C-ABI handle: model configuration, weights, and thin wrappers that turn
torch CUDA tensors into raw device pointers + the current CUDA stream.

PyTorch is used here only for device memory, streams and (elsewhere) torch.distributed.
"""
import ctypes as C
from typing import Dict, Optional, Sequence, Union

import torch

from . import _lib


class ZvError(RuntimeError):
    pass


def make_config(model_cfg: dict, vocab_size: int, pad_id: int = 0, distill: bool = False, dialog: bool = False,
                spk_a_id: int = 360, spk_b_id: int = 361, stereo: bool = False) -> _lib.ZvConfig:
    """model_cfg: the "model" section of model.json (= kwargs of ZipVoice.__init__, zipvoice.py:38-60)."""
    c = _lib.ZvConfig()
    ds = list(model_cfg["fm_decoder_downsampling_factor"])
    if len(ds) > _lib.ZV_MAX_STACKS:
        raise ValueError("too many encoder stacks")
    c.feat_dim = model_cfg["feat_dim"]
    c.fm_num_stacks = len(ds)
    for i, (d, n, k) in enumerate(zip(ds, model_cfg["fm_decoder_num_layers"],
                                      model_cfg["fm_decoder_cnn_module_kernel"])):
        c.fm_downsampling_factor[i], c.fm_num_layers[i], c.fm_cnn_module_kernel[i] = d, n, k
    c.fm_feedforward_dim = model_cfg["fm_decoder_feedforward_dim"]
    c.fm_num_heads = model_cfg["fm_decoder_num_heads"]
    c.fm_dim = model_cfg["fm_decoder_dim"]
    c.text_num_layers = model_cfg["text_encoder_num_layers"]
    c.text_feedforward_dim = model_cfg["text_encoder_feedforward_dim"]
    c.text_cnn_module_kernel = model_cfg["text_encoder_cnn_module_kernel"]
    c.text_num_heads = model_cfg["text_encoder_num_heads"]
    c.text_dim = model_cfg["text_encoder_dim"]
    for k in ("query_head_dim", "value_head_dim", "pos_head_dim", "pos_dim", "time_embed_dim", "text_embed_dim"):
        setattr(c, k, model_cfg[k])
    c.vocab_size, c.pad_id = vocab_size, pad_id
    c.use_guidance_scale_embed, c.dialog = int(distill), int(dialog)
    c.spk_a_id, c.spk_b_id = spk_a_id, spk_b_id
    c.stereo = int(stereo)
    return c


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class Engine:
    """One handle per process / GPU.  `state_dict` uses the reference's tensor names."""

    def __init__(self, model_cfg: dict, state_dict: Dict[str, torch.Tensor], vocab_size: int, pad_id: int = 0,
                 distill: bool = False, dialog: bool = False, spk_a_id: int = 360, spk_b_id: int = 361,
                 device: Union[str, torch.device] = "cuda:0", stereo: bool = False):
        if not torch.cuda.is_available():
            raise ZvError("zipvoice_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        self.device = torch.device(device)
        self.model_cfg, self.distill, self.dialog, self.stereo = dict(model_cfg), distill, dialog, stereo
        self.pad_id, self.feat_dim = pad_id, model_cfg["feat_dim"]
        self.x_dim = model_cfg["feat_dim"] * (2 if stereo else 1)  # width of x / speech_condition / velocity
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            cfg = make_config(model_cfg, vocab_size, pad_id, distill, dialog, spk_a_id, spk_b_id, stereo)
            self._check(self.lib.zv_create(C.byref(cfg), C.byref(self._h)))
            for name, t in state_dict.items():
                t = t.detach().to(dtype=torch.float32).contiguous()
                shape = (C.c_int64 * max(t.dim(), 1))(*t.shape)
                self._check(self.lib.zv_set_weight(self._h, name.encode(), _ptr(t), shape, t.dim(),
                                                   int(t.is_cuda)))
            self._check(self.lib.zv_finalize_weights(self._h))

    def _check(self, rc: int):
        if rc != 0:
            msg = self.lib.zv_last_error(self._h)
            raise ZvError(f"zipvoice_b200 error {rc}: {msg.decode() if msg else ''}")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self.lib.zv_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, key: str, value: int):
        self._check(self.lib.zv_set_option(self._h, key.encode(), int(value)))

    # ------------------------------------------------------------------ helpers
    def _select_stream(self, width: int):
        """TTSZipformerTwoStream.forward (zipformer_two_stream.py:238-244): the width of x picks the projection pair --
        2 * feat_dim: pair 0 (two channels), feat_dim: pair 1 (mono).  Single-stream models: nothing to choose."""
        if not self.stereo:
            return
        assert width in (2 * self.feat_dim, self.feat_dim), \
            f"{3 * width} not in ({5 * self.feat_dim}, {3 * self.feat_dim})"           # zipformer_two_stream.py:238
        stream = 0 if width == 2 * self.feat_dim else 1
        if stream != getattr(self, "_stream", 0):
            self.set_option("stream", stream)
            self._stream = stream
        self.x_dim = width

    def _prep(self, t: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
        return t.to(device=self.device, dtype=dtype).contiguous()

    def _guidance(self, guidance_scale, B: int):
        if torch.is_tensor(guidance_scale):
            g = guidance_scale.detach().float().reshape(-1).cpu().tolist()
        else:
            g = [float(guidance_scale)]
        assert len(g) in (1, B), "guidance_scale must be a float or a tensor of shape (batch, 1, 1)"
        return (C.c_float * len(g))(*g), len(g)

    # ------------------------------------------------------------------ hot path
    def velocity(self, t: float, x, text_condition, speech_condition, padding_mask=None, guidance_scale=0.0,
                 distill: Optional[bool] = None) -> torch.Tensor:
        """DiffusionModel.forward / DistillDiffusionModel.forward (solver.py:40-110,127-165)."""
        distill = self.distill if distill is None else distill
        self._select_stream(x.shape[-1])
        x, tc, sc = self._prep(x), self._prep(text_condition), self._prep(speech_condition)
        B, T, F = x.shape
        # zipformer.py:268 / zipformer_two_stream.py:240: the concatenated width must be the in_proj's
        assert F == self.x_dim and sc.shape == x.shape and tuple(tc.shape) == (B, T, self.feat_dim), \
            f"x/speech (B,T,{self.x_dim}) and text (B,T,{self.feat_dim}) expected, got {tuple(x.shape)} {tuple(sc.shape)} {tuple(tc.shape)}"
        m = None if padding_mask is None else self._prep(padding_mask, torch.uint8)
        g, glen = self._guidance(guidance_scale, B)
        out = torch.empty_like(x)
        with torch.cuda.device(self.device):
            self._check(self.lib.zv_velocity(self._h, int(distill), float(t), _ptr(x), _ptr(tc), _ptr(sc), _ptr(m), B,
                                             T, g, glen, _ptr(out), _stream(self.device)))
        return out

    def forward_fm_decoder(self, t, xt, text_condition, speech_condition, padding_mask=None,
                           guidance_scale=None) -> torch.Tensor:
        """ZipVoice.forward_fm_decoder (zipvoice.py:135-185): t 0-dim or (N, 1, 1) with one time step per row;
        guidance_scale None, 0-dim or (N, 1, 1) (embedded by the distilled model).  No CFG logic."""
        self._select_stream(xt.shape[-1])
        x, tc, sc = self._prep(xt), self._prep(text_condition), self._prep(speech_condition)
        N, T, F = x.shape
        assert F == self.x_dim and sc.shape == x.shape and tuple(tc.shape) == (N, T, self.feat_dim), \
            f"x/speech (N,T,{self.x_dim}) and text (N,T,{self.feat_dim}) expected"
        tv = torch.as_tensor(t).detach().float().reshape(-1).cpu().tolist()
        assert len(tv) in (1, N), "t must be 0-dim or (N, 1, 1)"                     # zipvoice.py:165-169
        if guidance_scale is None:
            gv = []
        else:
            gv = torch.as_tensor(guidance_scale).detach().float().reshape(-1).cpu().tolist()
            assert len(gv) in (1, N), "guidance_scale must be 0-dim or (N, 1, 1)"    # zipvoice.py:171-178
        m = None if padding_mask is None else self._prep(padding_mask, torch.uint8)
        out = torch.empty_like(x)
        with torch.cuda.device(self.device):
            self._check(self.lib.zv_forward_fm_decoder(
                self._h, (C.c_float * len(tv))(*tv), len(tv), _ptr(x), _ptr(tc), _ptr(sc), _ptr(m), N, T,
                (C.c_float * max(1, len(gv)))(*gv) if gv else None, len(gv), _ptr(out), _stream(self.device)))
        return out

    def sample(self, x, text_condition, speech_condition, padding_mask, num_step=10, guidance_scale=0.0,
               t_start=0.0, t_end=1.0, t_shift=1.0, distill: Optional[bool] = None) -> torch.Tensor:
        """EulerSolver.sample (solver.py:182-240); the input x is not modified."""
        distill = self.distill if distill is None else distill
        self._select_stream(x.shape[-1])
        x, tc, sc = self._prep(x), self._prep(text_condition), self._prep(speech_condition)
        B, T, F = x.shape
        # zipformer.py:268 / zipformer_two_stream.py:240: the concatenated width must be the in_proj's
        assert F == self.x_dim and sc.shape == x.shape and tuple(tc.shape) == (B, T, self.feat_dim), \
            f"x/speech (B,T,{self.x_dim}) and text (B,T,{self.feat_dim}) expected, got {tuple(x.shape)} {tuple(sc.shape)} {tuple(tc.shape)}"
        m = None if padding_mask is None else self._prep(padding_mask, torch.uint8)
        g, glen = self._guidance(guidance_scale, B)
        out = torch.empty_like(x)
        with torch.cuda.device(self.device):
            self._check(self.lib.zv_sample(self._h, int(distill), _ptr(x), _ptr(tc), _ptr(sc), _ptr(m), B, T,
                                           int(num_step), g, glen, float(t_start), float(t_end), float(t_shift),
                                           _ptr(out), _stream(self.device)))
        return out

    def text_embed(self, tokens_padded: torch.Tensor, tokens_lens: torch.Tensor) -> torch.Tensor:
        tp, tl = self._prep(tokens_padded, torch.int64), self._prep(tokens_lens, torch.int64)
        B, S = tp.shape
        out = torch.empty(B, S, self.feat_dim, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            self._check(self.lib.zv_text_embed(self._h, _ptr(tp), _ptr(tl), B, S, _ptr(out), _stream(self.device)))
        return out

    def text_condition(self, embed: torch.Tensor, tokens_lens: torch.Tensor, features_lens: torch.Tensor,
                       num_frames: int):
        emb = self._prep(embed)
        tl, fl = self._prep(tokens_lens, torch.int64), self._prep(features_lens, torch.int64)
        B, S, F = emb.shape
        out = torch.empty(B, num_frames, F, device=self.device, dtype=torch.float32)
        mask = torch.empty(B, num_frames, device=self.device, dtype=torch.uint8)
        with torch.cuda.device(self.device):
            self._check(self.lib.zv_text_condition(self._h, _ptr(emb), _ptr(tl), _ptr(fl), B, S, num_frames,
                                                   _ptr(out), _ptr(mask), _stream(self.device)))
        return out, mask.bool()

    # ------------------------------------------------------------------ audio front-end / vocoder
    def fbank(self, wav: torch.Tensor, lens: Optional[torch.Tensor] = None, kind: str = "vocos") -> torch.Tensor:
        """VocosFbank log-mel (feature.py:56-116), or with kind="bigvgan" the BigVGANFbank features (feature.py:133-204):
        wav (B, N) fp32 24 kHz mono -> (B, (N + 128) // 256, 100)."""
        assert kind in ("vocos", "bigvgan"), kind
        w = self._prep(wav)
        B, N = w.shape
        ln = None if lens is None else self._prep(lens, torch.int64)
        # torch.stft's reflect padding (feature.py:47-60 via torchaudio) needs more samples than the pad and raises
        # otherwise; the kernel would clamp silently, so short rows are rejected here (one host read of lens)
        pad = 512 if kind == "vocos" else 384
        shortest = N if ln is None else (int(ln.min()) if ln.numel() else N)
        if shortest <= pad:
            raise ValueError(f"fbank: every row needs more than {pad} samples (reflect padding), got {shortest}")
        T = (N + 128) // 256
        out = torch.empty(B, T, 100, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            fn = self.lib.zv_fbank if kind == "vocos" else self.lib.zv_fbank_bigvgan
            self._check(fn(self._h, _ptr(w), _ptr(ln), B, N, _ptr(out), _stream(self.device)))
        return out

    def resample(self, wav: torch.Tensor, orig_freq: int, new_freq: int,
                 lens: Optional[torch.Tensor] = None) -> torch.Tensor:
        """torchaudio.transforms.Resample(orig_freq, new_freq) (bin/infer_zipvoice.py:334-339): (B, N) -> (B, n_out),
        n_out = ceil(new_freq * N / orig_freq); with ``lens`` every row is resampled as if it ended at lens[b]."""
        w = self._prep(wav)
        B, N = w.shape
        ln = None if lens is None else self._prep(lens, torch.int64)
        n_out = (N * new_freq + orig_freq - 1) // orig_freq
        out = torch.empty(B, n_out, device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            self._check(self.lib.zv_resample(self._h, _ptr(w), _ptr(ln), B, N, int(orig_freq), int(new_freq),
                                             _ptr(out), n_out, _stream(self.device)))
        return out

    RESAMPLE_STATIC_MAX_FLOATS = 65536  # ZV_RESAMPLE_STATIC_MAX_FLOATS (include/zipvoice_b200.h)

    def resample_static(self, wav: torch.Tensor, orig_freq: int, new_freq: int, out: torch.Tensor,
                        lens: Optional[torch.Tensor] = None) -> torch.Tensor:
        """zv_resample_static: ``resample`` for callers that must not allocate (no counterpart in the reference).
        The CALLER owns every buffer: ``wav`` (B, N) fp32 and ``out``
        (B, ceil(new_freq * N / orig_freq)) fp32, contiguous, on this engine's device (``lens`` int64 likewise) --
        nothing is converted, copied or allocated here or in the library.  The filter bank lives in a 256 KB static
        device array shared by all calls on the GPU: rate pairs whose bank exceeds RESAMPLE_STATIC_MAX_FLOATS raise,
        and calls that overlap on the GPU (different streams) race."""
        for t, dt in ((wav, torch.float32), (out, torch.float32)) + (() if lens is None else ((lens, torch.int64),)):
            assert t.is_cuda and t.device == self.device and t.dtype == dt and t.is_contiguous(), \
                "resample_static takes caller-owned contiguous tensors on the engine's device"
        B, N = wav.shape
        n_out = (N * new_freq + orig_freq - 1) // orig_freq
        assert tuple(out.shape) == (B, n_out), f"out must be ({B}, {n_out})"
        with torch.cuda.device(self.device):
            self._check(self.lib.zv_resample_static(self._h, _ptr(wav), _ptr(lens), B, N, int(orig_freq),
                                                    int(new_freq), _ptr(out), n_out, _stream(self.device)))
        return out

    def vocos_decode(self, mel: torch.Tensor, lens: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Vocos.decode: mel (B, 100, T) -> waveform (B, 256 (T - 1)); needs "vocoder.*" weights."""
        m = self._prep(mel)
        B, C, T = m.shape
        assert C == 100
        ln = None if lens is None else self._prep(lens, torch.int64)
        out = torch.empty(B, 256 * (T - 1), device=self.device, dtype=torch.float32)
        with torch.cuda.device(self.device):
            self._check(self.lib.zv_vocos_decode(self._h, _ptr(m), _ptr(ln), B, T, _ptr(out), _stream(self.device)))
        return out

    def prof_enable(self, on: bool = True):
        self._check(self.lib.zv_prof_enable(self._h, int(on)))

    def prof_read(self):
        """-> list of dicts {name, tensor, ms, work, launches} per kernel class (device is synchronised)."""
        n = self.lib.zv_prof_num_classes()
        ms, work, cnt = (C.c_double * n)(), (C.c_double * n)(), (C.c_longlong * n)()
        self._check(self.lib.zv_prof_read(self._h, ms, work, cnt))
        return [dict(name=self.lib.zv_prof_class_name(i).decode(), tensor=bool(self.lib.zv_prof_class_is_tensor(i)),
                     ms=ms[i], work=work[i], launches=int(cnt[i])) for i in range(n) if cnt[i]]

    def launch_count_reset(self):
        self.lib.zv_launch_count_reset()

    def launch_count(self) -> int:
        return int(self.lib.zv_launch_count())

"""Drop-in replacements for the reference solver objects (zipvoice/models/modules/solver.py).

``model.solver = B200EulerSolver(engine)`` is the plug-in point (zipvoice.py:133,460-468):
``ZipVoice.sample()`` / ``sample_intermediate()`` call ``self.solver.sample(x=..., text_condition=...,
speech_condition=..., padding_mask=..., num_step=..., guidance_scale=..., t_shift=...)`` by keyword.
Same argument names, defaults and assertions as the reference; the whole loop (CFG batch doubling,
fm_decoder, CFG combine, Euler update) runs inside one C-ABI call with no host synchronisation.
"""
from typing import Optional, Union

import torch

from . import _lib
from .engine import Engine


def get_time_steps(t_start: float = 0.0, t_end: float = 1.0, num_step: int = 10, t_shift: float = 1.0,
                   device: torch.device = torch.device("cpu")) -> torch.Tensor:
    """solver.py:256-281, evaluated by the library's own host routine (bit-identical to torch fp32)."""
    import ctypes as C
    buf = (C.c_float * (num_step + 1))()
    _lib.load().zv_time_steps(float(t_start), float(t_end), int(num_step), float(t_shift), buf)
    return torch.tensor(list(buf), dtype=torch.float32, device=device)


class B200DiffusionModel:
    """DiffusionModel.forward (solver.py:40-110): one guided velocity evaluation."""

    distill = False

    def __init__(self, engine: Engine):
        self.engine = engine

    def __call__(self, t: torch.Tensor, x: torch.Tensor, text_condition: torch.Tensor,
                 speech_condition: torch.Tensor, padding_mask: Optional[torch.Tensor] = None,
                 guidance_scale: Union[float, torch.Tensor] = 0.0, **kwargs) -> torch.Tensor:
        t = torch.as_tensor(t)
        assert t.dim() == 0, "the B200 path takes one scalar time step per call (solver.py:81)"
        return self.engine.velocity(float(t), x, text_condition, speech_condition, padding_mask, guidance_scale,
                                    distill=self.distill)

    forward = __call__


class B200DistillDiffusionModel(B200DiffusionModel):
    """DistillDiffusionModel.forward (solver.py:127-165): guidance enters as an embedding, no doubling."""

    distill = True


class B200EulerSolver:
    def __init__(self, engine: Engine, func_name: str = "forward_fm_decoder"):
        assert func_name == "forward_fm_decoder", "only the fm_decoder velocity function is accelerated"
        self.engine = engine
        self.model = B200DiffusionModel(engine)

    def sample(self, x: torch.Tensor, text_condition: torch.Tensor, speech_condition: torch.Tensor,
               padding_mask: torch.Tensor, num_step: int = 10, guidance_scale: Union[float, torch.Tensor] = 0.0,
               t_start: float = 0.0, t_end: float = 1.0, t_shift: float = 1.0, **kwargs) -> torch.Tensor:
        assert isinstance(t_start, float) and isinstance(t_end, float)  # solver.py:219
        return self.engine.sample(x, text_condition, speech_condition, padding_mask, num_step=num_step,
                                  guidance_scale=guidance_scale, t_start=t_start, t_end=t_end, t_shift=t_shift,
                                  distill=self.model.distill)


class B200DistillEulerSolver(B200EulerSolver):
    def __init__(self, engine: Engine, func_name: str = "forward_fm_decoder"):
        super().__init__(engine, func_name)
        self.model = B200DistillDiffusionModel(engine)

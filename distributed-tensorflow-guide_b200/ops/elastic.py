"""ctypes wrappers for ``csrc/elastic_kernels.cu`` (DynSGD / AEASGD / AEAMSGD data plane)."""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _check, _stream, lib

_declared = False


def _l():
    global _declared
    l = lib()
    if not _declared:
        vp, i64, f32 = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_float
        l.dtg_easgd_elastic.argtypes = [vp, vp, vp, vp, f32, f32, i64, vp]
        l.dtg_momentum_step.argtypes = [vp, vp, vp, vp, f32, f32, i64, vp]
        l.dtg_push_axpy_dyn.argtypes = [vp, vp, f32, vp, vp, vp, i64, vp]
        for n in ("dtg_easgd_elastic", "dtg_momentum_step", "dtg_push_axpy_dyn"):
            getattr(l, n).restype = ctypes.c_int
        _declared = True
    return l


def easgd_elastic(centre: torch.Tensor, x: torch.Tensor, x_bf16: Optional[torch.Tensor], alpha: float,
                  velocity: Optional[torch.Tensor] = None, delta: float = 0.0):
    """``e = alpha (x - centre); x -= e; centre += e`` (centre may be peer memory); the bf16
    mirror is refreshed with ``x`` (or the look-ahead ``x + delta v`` when ``velocity`` is given)."""
    _check(_l().dtg_easgd_elastic(centre.data_ptr(), x.data_ptr(),
                                  None if x_bf16 is None else x_bf16.data_ptr(),
                                  None if velocity is None else velocity.data_ptr(), float(delta),
                                  float(alpha), x.numel(), _stream()), "dtg_easgd_elastic")


def momentum_step(x, v, g, x_bf16, lr: float, delta: float):
    """``v = delta v - lr g; x += v``; the bf16 mirror holds the Nesterov look-ahead ``x + delta v``."""
    _check(_l().dtg_momentum_step(x.data_ptr(), v.data_ptr(), g.data_ptr(),
                                  None if x_bf16 is None else x_bf16.data_ptr(), float(lr),
                                  float(delta), x.numel(), _stream()), "dtg_momentum_step")


def push_axpy_dyn(w_ps, g, alpha: float, step_word, seen, staleness_out=None):
    """``w_ps += alpha / staleness * g`` with ``staleness = step_word - seen + 1`` read on device."""
    _check(_l().dtg_push_axpy_dyn(w_ps.data_ptr(), g.data_ptr(), float(alpha), step_word.data_ptr(),
                                  seen.data_ptr(),
                                  None if staleness_out is None else staleness_out.data_ptr(),
                                  g.numel(), _stream()), "dtg_push_axpy_dyn")

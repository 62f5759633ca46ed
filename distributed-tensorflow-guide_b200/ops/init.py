"""ctypes wrapper for ``csrc/init_kernels.cu``: variable initialisers on the device, and
``device_init`` which fills a native model's flat fp32 buffer from its ``ParamSpec.dist`` tags.

Reference sites: ``tf.truncated_normal([2])`` (/root/reference/Non-Distributed_Setup.py:13-14),
``tf.constant(0.)`` (/root/reference/DOWNPOUR/DOWNPOUR.py:46-48); SURVEY 2.5 K16, section 8 W15.
"""
from __future__ import annotations

import ctypes
import zlib
from typing import Optional

import torch

from . import _check, _stream, lib

__all__ = ["NORMAL", "TRUNC_NORMAL", "UNIFORM", "init_random_", "device_init"]

NORMAL, TRUNC_NORMAL, UNIFORM = 0, 1, 2
_KINDS = {"normal": NORMAL, "trunc_normal": TRUNC_NORMAL, "uniform": UNIFORM}
_declared = False


def _l():
    global _declared
    l = lib()
    if not _declared:
        vp, i64, f32 = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_float
        l.dtg_init_random.argtypes = [vp, i64, i64, i64, i64, ctypes.c_int, f32, f32,
                                      ctypes.c_ulonglong, ctypes.c_uint, vp]
        l.dtg_init_random.restype = ctypes.c_int
        _declared = True
    return l


def _fill(t: torch.Tensor, first: int, kind: int, a: float, b: float, seed: int, stream_id: int):
    """``t``: fp32 view, contiguous, or 2-D with unit inner stride; higher-rank strided views
    (padding in a middle dimension) are walked along dim 0."""
    if t.numel() == 0:
        return
    if t.is_contiguous():
        rows, cols, ld = 1, t.numel(), t.numel()
    elif t.dim() == 2 and t.stride(1) == 1:
        rows, cols, ld = t.shape[0], t.shape[1], t.stride(0)
    else:
        per = t[0].numel()
        for i in range(t.shape[0]):
            _fill(t[i], first + i * per, kind, a, b, seed, stream_id)
        return
    _check(_l().dtg_init_random(t.data_ptr(), rows, cols, ld, first, kind, float(a), float(b),
                                seed & 0xFFFFFFFFFFFFFFFF, stream_id & 0xFFFFFFFF, _stream()),
           "dtg_init_random")


def init_random_(t: torch.Tensor, kind, a: float = 0.0, b: float = 1.0, seed: int = 0,
                 stream_id: int = 0) -> torch.Tensor:
    """In place: ``kind`` NORMAL ``a + b z`` / TRUNC_NORMAL (|z| <= 2, redrawn) / UNIFORM ``[a, b)``.
    The value of element ``e`` depends only on ``(seed, stream_id, e)``."""
    if t.dtype != torch.float32 or not t.is_cuda:
        raise ValueError("init_random_ needs a CUDA fp32 tensor")
    _fill(t, 0, _KINDS.get(kind, kind), a, b, seed, stream_id)
    return t


def device_init(model, layout, p32: torch.Tensor, seed: int = 0,
                host_fallback: Optional[bool] = True) -> int:
    """Initialise every parameter of ``model`` directly in the flat fp32 buffer ``p32``.

    ``ParamSpec.dist`` says how: ``("normal", std)``, ``("trunc_normal", std)``,
    ``("uniform", bound)`` (in ``[-bound, bound)``) or ``("constant", value)``; the random stream
    of a tensor is a hash of its name, so the result does not depend on creation order or on
    which ps shard the tensor landed in.  Padding of the stored layout stays zero.  Specs
    without a tag are drawn on the host (``host_fallback``) or refused.  Returns the number of
    tensors initialised on the device."""
    p32.zero_()
    n_dev = 0
    gen = None
    for s in model.param_specs():
        if s.name not in layout.entries:
            continue
        v = layout.logical(p32, s.name)
        dist = getattr(s, "dist", None)
        if dist is None:
            if not host_fallback:
                raise ValueError("no device initialiser for %s" % s.name)
            if gen is None:
                gen = torch.Generator().manual_seed(seed)
            v.copy_(s.init(s.shape, gen).to(p32.device))
            continue
        kind, x = dist
        sid = zlib.crc32(s.name.encode())
        if kind == "constant":
            if x != 0.0:
                v.fill_(x)
        elif kind == "uniform":
            _fill(v, 0, UNIFORM, -x, x, seed, sid)
        else:
            _fill(v, 0, _KINDS[kind], 0.0, x, seed, sid)
        n_dev += 1
    return n_dev

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

__all__ = ["NORMAL", "TRUNC_NORMAL", "UNIFORM", "init_random_", "device_init", "host_random"]

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


def _philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10 on numpy uint64 arrays holding 32-bit words."""
    import numpy as np
    M0, M1, MASK = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), np.uint64(0xFFFFFFFF)
    k0 = np.uint64(k0)
    k1 = np.uint64(k1)
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        c0, c1, c2, c3 = ((p1 >> np.uint64(32)) ^ c1 ^ k0) & MASK, p1 & MASK, \
            ((p0 >> np.uint64(32)) ^ c3 ^ k1) & MASK, p0 & MASK
        k0 = (k0 + np.uint64(0x9E3779B9)) & MASK
        k1 = (k1 + np.uint64(0xBB67AE85)) & MASK
    return c0, c1, c2, c3


def host_random(n: int, kind, a: float = 0.0, b: float = 1.0, seed: int = 0, stream_id: int = 0,
                first: int = 0) -> torch.Tensor:
    """The kernel's stream on the host (numpy): element ``first + i`` of stream ``stream_id``
    under ``seed``, for ``i < n``.  Uniform values are bit-identical to the device's; normal
    ones agree to the accuracy of the device's fast ``log`` / ``cos`` (~1e-6 relative).  Lets
    CPU code start from the parameters ``device_init`` would produce, and checks the kernel."""
    import numpy as np
    kind = _KINDS.get(kind, kind)
    e = np.arange(first, first + n, dtype=np.uint64)
    lo, hi = e & np.uint64(0xFFFFFFFF), e >> np.uint64(32)
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    sid = np.full(n, stream_id & 0xFFFFFFFF, dtype=np.uint64)

    def u01(x):                                    # (0, 1], as on the device
        return ((x >> np.uint64(8)).astype(np.float32) + np.float32(1.0)) * np.float32(1.0 / 16777216.0)

    def normal(x, y):
        r = np.sqrt(np.float32(-2.0) * np.log(u01(x)))
        return (r * np.cos(np.float32(6.283185307179586) * u01(y))).astype(np.float32)
    if kind == UNIFORM:
        r0, _, _, _ = _philox4x32_10(lo, hi, np.zeros(n, np.uint64), sid, k0, k1)
        v = np.float32(a) + np.float32(b - a) * ((r0 >> np.uint64(8)).astype(np.float32)
                                                 * np.float32(1.0 / 16777216.0))
        return torch.from_numpy(v.astype(np.float32))
    z = np.zeros(n, np.float32)
    todo = np.ones(n, bool)
    for attempt in range(16):
        idx = np.nonzero(todo)[0]
        if idx.size == 0:
            break
        r0, r1, r2, r3 = _philox4x32_10(lo[idx], hi[idx], np.full(idx.size, attempt, np.uint64),
                                        sid[idx], k0, k1)
        z1 = normal(r0, r1)
        if kind != TRUNC_NORMAL:
            z[idx] = z1
            todo[idx] = False
            break
        ok1 = np.abs(z1) <= 2.0
        z2 = normal(r2, r3)
        ok2 = ~ok1 & (np.abs(z2) <= 2.0)
        z[idx] = np.where(ok1, z1, z2)             # z2 stays as the last draw if both fail
        todo[idx] = ~(ok1 | ok2)
    if kind == TRUNC_NORMAL:
        z = np.clip(z, -2.0, 2.0)
    return torch.from_numpy((np.float32(a) + np.float32(b) * z).astype(np.float32))


def _fill(t: torch.Tensor, first: int, kind: int, a: float, b: float, seed: int, stream_id: int):
    """``t``: fp32 view, contiguous, or 2-D with unit inner stride; higher-rank strided views
    (padding in a middle dimension) are walked along dim 0."""
    if t.numel() == 0:
        return
    if not t.is_cuda:                              # same stream, drawn on the host
        t.copy_(host_random(t.numel(), kind, a, b, seed, stream_id, first).view(t.shape))
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
    if t.dtype != torch.float32:
        raise ValueError("init_random_ needs an fp32 tensor (CUDA: the kernel; CPU: host_random)")
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

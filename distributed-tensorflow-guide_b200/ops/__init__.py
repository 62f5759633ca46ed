"""sm_100a kernels, reached through the C ABI of ``libdtg_b200.so`` (ctypes).

The library is built in-tree by ``ops/build.py`` (``__graft_entry__.build()``).
On a machine with a GPU every op here *requires* the library: there is no
silent PyTorch fallback -- a missing or stale ``.so`` raises ``NativeLibraryError``.

GEMM operand conventions (bf16 in, fp32 accumulate, tcgen05 + TMEM + TMA):

* ``a_mn=False``: ``A`` is ``[M, K]`` row-major (K-major).  ``a_mn=True``: ``A`` is given
  as ``[K, M]`` row-major (MN-major) -- e.g. ``dY`` used as ``dY^T`` in a wgrad.
* ``b_mn=False``: ``B`` is ``[N, K]`` row-major (``torch.nn.Linear`` weight layout).
  ``b_mn=True``: ``B`` is ``[K, N]`` row-major.
* result ``D[M, N] = alpha * A.B (+ bias) -> act -> * mask' (+ residual)`` written as
  bf16, fp32 store, fp32 ``+=`` or fp32 ``red.add`` (lock-free push; the target may
  be a peer GPU's memory).
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

__all__ = ["lib", "available", "NativeLibraryError", "gemm", "linear_forward", "pull_convert",
           "push_axpy", "push_adagrad", "local_step", "softmax_xent", "bias_grad",
           "cast_f32_bf16", "counter_add", "EPI_BF16", "EPI_F32", "EPI_F32_ACC", "EPI_F32_RED",
           "ACT_NONE", "ACT_RELU", "ACT_GELU", "MASK_NONE", "MASK_RELU", "MASK_GELU",
           "launch_count", "reset_launch_count"]

EPI_BF16, EPI_F32, EPI_F32_ACC, EPI_F32_RED = 0, 1, 2, 3
ACT_NONE, ACT_RELU, ACT_GELU = 0, 1, 2
MASK_NONE, MASK_RELU, MASK_GELU = 0, 1, 2

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libdtg_b200.so")
_lib = None
_launches = 0


class NativeLibraryError(RuntimeError):
    pass


def _declare(l):
    vp, i32, i64, f32, u64 = (ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float,
                              ctypes.c_ulonglong)
    l.dtg_last_error.restype = ctypes.c_char_p
    l.dtg_version.restype = i32
    l.dtg_gemm_bf16.argtypes = [vp, vp, i32, i32, i32, i64, i64, i32, i32, vp, vp, vp, vp, vp,
                                i64, i64, f32, i32, i32, i32, i32, vp]
    l.dtg_pull_convert.argtypes = [vp, vp, vp, i64, vp]
    l.dtg_push_axpy.argtypes = [vp, vp, f32, i64, vp]
    l.dtg_push_adagrad.argtypes = [vp, vp, vp, f32, f32, i64, vp]
    l.dtg_local_step.argtypes = [vp, vp, vp, vp, vp, f32, i32, i64, vp]
    l.dtg_softmax_xent.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, f32, vp]
    l.dtg_bias_grad.argtypes = [vp, vp, i32, i32, i64, f32, vp]
    l.dtg_cast_f32_bf16.argtypes = [vp, vp, i64, vp]
    l.dtg_counter_add.argtypes = [vp, u64, vp]
    l.dtg_sync_rs_update_ag.argtypes = [vp, i32, vp]
    l.dtg_sync_params_size.restype = i32
    l.dtg_fabric_barrier.argtypes = [vp, i32, i32, i32, ctypes.c_uint, i64, vp, vp]
    l.dtg_ipc_handle_size.restype = i32
    l.dtg_ipc_alloc.argtypes = [ctypes.c_size_t, ctypes.POINTER(vp), vp]
    l.dtg_ipc_open.argtypes = [vp, ctypes.POINTER(vp)]
    l.dtg_ipc_close.argtypes = [vp]
    l.dtg_free.argtypes = [vp]
    l.dtg_enable_peer_access.argtypes = [i32, i32]
    l.dtg_memset_async.argtypes = [vp, i32, ctypes.c_size_t, vp]
    l.dtg_memcpy_async.argtypes = [vp, vp, ctypes.c_size_t, vp]
    for name in ("dtg_gemm_bf16", "dtg_pull_convert", "dtg_push_axpy", "dtg_push_adagrad",
                 "dtg_local_step", "dtg_softmax_xent", "dtg_bias_grad", "dtg_cast_f32_bf16",
                 "dtg_counter_add", "dtg_sync_rs_update_ag", "dtg_fabric_barrier",
                 "dtg_ipc_alloc", "dtg_ipc_open", "dtg_ipc_close", "dtg_free",
                 "dtg_enable_peer_access", "dtg_memset_async", "dtg_memcpy_async"):
        getattr(l, name).restype = i32


def lib():
    """The loaded shared library (loads on first use)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise NativeLibraryError(
                "%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a) before using dtg_b200.ops" % _LIB_PATH)
        try:
            l = ctypes.CDLL(_LIB_PATH)
            _declare(l)
        except (OSError, AttributeError) as e:
            raise NativeLibraryError("cannot load %s: %s" % (_LIB_PATH, e))
        _lib = l
    return _lib


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def _check(rc: int, what: str):
    global _launches
    if rc != 0:
        raise NativeLibraryError("%s failed (status %d): %s" % (
            what, rc, lib().dtg_last_error().decode(errors="replace")))
    _launches += 1


def launch_count() -> int:
    """Kernels launched through this module since the last reset (bench evidence)."""
    return _launches


def reset_launch_count() -> None:
    global _launches
    _launches = 0


def _stream(t: Optional[torch.Tensor] = None) -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _need(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise ValueError("%s must be a CUDA tensor" % name)
    if t.dtype != dtype:
        raise TypeError("%s must be %s, got %s" % (name, dtype, t.dtype))


# --------------------------------------------------------------------------- #
# GEMM
# --------------------------------------------------------------------------- #
def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, *, a_mn: bool = False,
         b_mn: bool = False, alpha: float = 1.0, bias: Optional[torch.Tensor] = None,
         act: int = ACT_NONE, mask: Optional[torch.Tensor] = None, mask_mode: int = MASK_NONE,
         residual: Optional[torch.Tensor] = None, out2: Optional[torch.Tensor] = None,
         mode: Optional[int] = None, force_bn: int = 0) -> torch.Tensor:
    """``out[M,N] (op)= epilogue(alpha * A.B)`` on the tcgen05 tensor cores.

    ``a``/``b`` are 2-D bf16 with unit inner stride (row pitch may exceed the width).
    ``mode`` defaults to ``EPI_BF16`` for a bf16 ``out`` and ``EPI_F32`` for fp32.
    """
    _need(a, torch.bfloat16, "a")
    _need(b, torch.bfloat16, "b")
    if a.dim() != 2 or b.dim() != 2 or a.stride(1) != 1 or b.stride(1) != 1:
        raise ValueError("gemm operands must be 2-D with unit inner stride")
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    if K != Kb:
        raise ValueError("inner dimensions differ: %d vs %d" % (K, Kb))
    if out.dim() != 2 or out.shape[0] != M or out.shape[1] < N or out.stride(1) != 1:
        raise ValueError("out must be [M, >=N] with unit inner stride")
    if mode is None:
        mode = EPI_BF16 if out.dtype == torch.bfloat16 else EPI_F32
    if mode == EPI_BF16:
        _need(out, torch.bfloat16, "out")
    else:
        _need(out, torch.float32, "out")
    if bias is not None:
        _need(bias, torch.float32, "bias")
    if mask is not None:
        _need(mask, torch.bfloat16, "mask")
        if mask_mode == MASK_NONE:
            mask_mode = MASK_RELU
    if residual is not None:
        _need(residual, torch.bfloat16, "residual")
        if residual.stride(0) != out.stride(0):
            raise ValueError("residual must share out's row pitch")
    if out2 is not None:
        _need(out2, torch.bfloat16, "out2")
        if out2.stride(0) != out.stride(0):
            raise ValueError("out2 must share out's row pitch")
    rc = lib().dtg_gemm_bf16(
        a.data_ptr(), b.data_ptr(), M, N, K, a.stride(0), b.stride(0), int(a_mn), int(b_mn),
        out.data_ptr(), _ptr(out2), _ptr(bias), _ptr(mask), _ptr(residual), out.stride(0),
        mask.stride(0) if mask is not None else 0, float(alpha), int(mode), int(act),
        int(mask_mode), int(force_bn), _stream())
    _check(rc, "dtg_gemm_bf16")
    return out


def linear_forward(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None,
                   act: int = ACT_NONE, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``act(x @ w.T + bias)``; ``x [B, in]``, ``w [out, in]`` bf16."""
    if out is None:
        out = torch.empty(x.shape[0], w.shape[0], device=x.device, dtype=torch.bfloat16)
    return gemm(x, w, out, bias=bias, act=act)


# --------------------------------------------------------------------------- #
# parameter-server data plane
# --------------------------------------------------------------------------- #
def pull_convert(src_f32: torch.Tensor, dst_bf16: Optional[torch.Tensor] = None,
                 dst_f32: Optional[torch.Tensor] = None) -> None:
    """PULL: ``dst_bf16 <- bf16(src)``, ``dst_f32 <- src``; ``src`` may be peer memory."""
    n = src_f32.numel()
    _check(lib().dtg_pull_convert(src_f32.data_ptr(), _ptr(dst_bf16), _ptr(dst_f32), n, _stream()),
           "dtg_pull_convert")


def push_axpy(w_ps: torch.Tensor, g: torch.Tensor, alpha: float) -> None:
    """PUSH (SGD on the ps): ``w_ps += alpha * g`` with ``red.global.add.v4.f32``."""
    _check(lib().dtg_push_axpy(w_ps.data_ptr(), g.data_ptr(), float(alpha), g.numel(), _stream()),
           "dtg_push_axpy")


def push_adagrad(w_ps: torch.Tensor, acc_ps: torch.Tensor, g: torch.Tensor, lr: float,
                 gscale: float = 1.0) -> None:
    """PUSH (Adagrad on the ps): ``acc += g^2`` (atomic) then ``w -= lr*g*rsqrt(acc)``."""
    _check(lib().dtg_push_adagrad(w_ps.data_ptr(), acc_ps.data_ptr(), g.data_ptr(), float(lr),
                                  float(gscale), g.numel(), _stream()), "dtg_push_adagrad")


def local_step(p: torch.Tensor, g: torch.Tensor, lr: float, *, p_bf16: Optional[torch.Tensor] = None,
               slot: Optional[torch.Tensor] = None, window: Optional[torch.Tensor] = None,
               adagrad: bool = False) -> None:
    """Fused worker-local optimizer step + window accumulate over flat buffers."""
    _check(lib().dtg_local_step(p.data_ptr(), _ptr(p_bf16), _ptr(slot), _ptr(window), g.data_ptr(),
                                float(lr), int(adagrad), p.numel(), _stream()), "dtg_local_step")


def softmax_xent(logits: torch.Tensor, labels: torch.Tensor, dlogits: Optional[torch.Tensor],
                 loss_sum: torch.Tensor, num_classes: int, scale: float) -> None:
    """``loss_sum += scale * sum(-log softmax[label])``; ``dlogits = scale*(p - onehot)``."""
    _need(logits, torch.float32, "logits")
    _need(labels, torch.int64, "labels")
    B = logits.shape[0]
    _check(lib().dtg_softmax_xent(logits.data_ptr(), labels.data_ptr(), _ptr(dlogits),
                                  loss_sum.data_ptr(), B, num_classes, logits.stride(0),
                                  dlogits.stride(0) if dlogits is not None else 0, float(scale),
                                  _stream()), "dtg_softmax_xent")


def bias_grad(dy: torch.Tensor, target: torch.Tensor, alpha: float = 1.0,
              n: Optional[int] = None) -> None:
    """``target[n] += alpha * sum_m dy[m, n]`` (reduction add; target may be peer memory)."""
    _need(dy, torch.bfloat16, "dy")
    M = dy.shape[0]
    N = n if n is not None else dy.shape[1]
    _check(lib().dtg_bias_grad(dy.data_ptr(), target.data_ptr(), M, N, dy.stride(0), float(alpha),
                               _stream()), "dtg_bias_grad")


def cast_f32_bf16(src: torch.Tensor, dst: torch.Tensor) -> None:
    _check(lib().dtg_cast_f32_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _stream()),
           "dtg_cast_f32_bf16")


def counter_add(counter: torch.Tensor, delta: int = 1) -> None:
    """``global_step += delta`` on a (possibly peer-resident) int64 word."""
    _check(lib().dtg_counter_add(counter.data_ptr(), int(delta), _stream()), "dtg_counter_add")

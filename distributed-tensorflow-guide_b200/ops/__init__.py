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
           "push_axpy", "push_adagrad", "local_step", "softmax_xent", "softmax_xent_bf16_", "ZeroMany", "pull_ranges", "bias_grad",
           "cast_f32_bf16", "counter_add", "EPI_BF16", "EPI_F32", "EPI_F32_ACC", "EPI_F32_RED",
           "EPI_F32_RED_COAL",
           "ACT_NONE", "ACT_RELU", "ACT_GELU", "MASK_NONE", "MASK_RELU", "MASK_GELU",
           "launch_count", "reset_launch_count", "gemm_batched", "layernorm_fwd", "layernorm_bwd",
           "embedding_fwd", "embedding_bwd", "attn_softmax_fwd", "attn_softmax_bwd", "add_bf16"]

EPI_BF16, EPI_F32, EPI_F32_ACC, EPI_F32_RED = 0, 1, 2, 3
EPI_F32_RED_COAL = 4   # reduction-add staged for 128 B NVLink payloads (remote ps shard)
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
    l.dtg_gemm_bf16_batched.argtypes = [vp, vp, i32, i32, i32, i64, i64, i32, i32, vp, i64, f32,
                                        i32, i32, i32, i64, i64, i64, i64, i64, i64, i32, vp]
    l.dtg_layernorm_fwd.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, f32, vp]
    l.dtg_layernorm_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, f32, vp]
    l.dtg_embedding_fwd.argtypes = [vp, vp, vp, vp, i32, i32, i32, vp]
    l.dtg_embedding_bwd.argtypes = [vp, vp, vp, vp, i32, i32, i32, f32, vp]
    l.dtg_attn_softmax_fwd.argtypes = [vp, vp, i64, i32, f32, i32, vp]
    l.dtg_attn_softmax_bwd.argtypes = [vp, vp, vp, i64, i32, f32, vp]
    l.dtg_add_bf16.argtypes = [vp, vp, vp, i64, vp]
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
                 "dtg_gemm_bf16_batched", "dtg_layernorm_fwd", "dtg_layernorm_bwd",
                 "dtg_embedding_fwd", "dtg_embedding_bwd", "dtg_attn_softmax_fwd",
                 "dtg_attn_softmax_bwd", "dtg_add_bf16",
                 "dtg_ipc_alloc", "dtg_ipc_open", "dtg_ipc_close", "dtg_free",
                 "dtg_enable_peer_access", "dtg_memset_async", "dtg_memcpy_async"):
        getattr(l, name).restype = i32


def _ensure_current():
    """A library built from other sources than the ones next to it must never be the one that
    runs (ops/build.py stamps the .so with a content hash of csrc/).  If it is stale and nvcc
    is here, rebuild under a file lock (N ranks start at once); if it cannot be rebuilt, say so."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("dtg_b200_ops_build", os.path.join(_HERE, "build.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    if b.is_current() or os.environ.get("DTG_ALLOW_STALE_LIB") == "1":
        return
    import fcntl
    with open(os.path.join(_HERE, ".build.lock"), "w") as lk:
        fcntl.flock(lk, fcntl.LOCK_EX)
        try:
            if not b.is_current():
                try:
                    b.build()
                except RuntimeError as e:
                    raise NativeLibraryError("libdtg_b200.so does not match ops/csrc and cannot be "
                                             "rebuilt here: %s" % e)
        finally:
            fcntl.flock(lk, fcntl.LOCK_UN)


def lib():
    """The loaded shared library (loads on first use)."""
    global _lib
    if _lib is None:
        _ensure_current()
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
    elif mode < 100:                     # >= 100: timing experiments that write nothing
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


def tall_schedule_check(M: int, N: int, clusters: int = 74, small_cost: float = 0.62,
                        group: int = 8) -> dict:
    """Host replay of the tall-pair GEMM's static tile schedule for an ``M x N`` output on
    ``clusters`` CTA pairs (``dtg_tall_schedule_check``; no GPU needed): raises if any 256 x 256
    output block would be produced zero or several times, else returns the plan."""
    l = lib()
    l.dtg_tall_schedule_check.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float,
                                          ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    l.dtg_tall_schedule_check.restype = ctypes.c_int
    info = (ctypes.c_int * 6)()
    _check(l.dtg_tall_schedule_check(int(M), int(N), int(clusters), float(small_cost), int(group),
                                     info), "dtg_tall_schedule_check")
    return {"big_mblocks": info[0], "big_tiles": info[1], "small_tiles": info[2],
            "extra_rounds": info[3], "clusters": info[4], "worst_load": info[5] / 1000.0}


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


class ZeroMany:
    """``ZeroMany(tensors)()`` zeroes all of them (fp32, contiguous) with one kernel launch; the
    address table is built once and kept on the device."""

    def __init__(self, tensors):
        ts = [t for t in tensors if t.numel() > 0]
        for t in ts:
            _need(t, torch.float32, "tensor")
            if not t.is_contiguous():
                raise ValueError("ZeroMany needs contiguous tensors")
        self._keep = ts
        self.count = len(ts)
        self.longest = max([t.numel() for t in ts], default=0)
        flat = []
        for t in ts:
            flat += [t.data_ptr(), t.numel()]
        self.table = (torch.tensor(flat, dtype=torch.int64, device=ts[0].device) if ts else None)

    def __call__(self) -> None:
        if self.count == 0:
            return
        l = lib()
        l.dtg_zero_many.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p]
        l.dtg_zero_many.restype = ctypes.c_int
        _check(l.dtg_zero_many(self.table.data_ptr(), self.count, self.longest, _stream()),
               "dtg_zero_many")


def pull_ranges(src_f32: torch.Tensor, dst_f32: torch.Tensor, ranges) -> None:
    """``dst[lo:hi] = src[lo:hi]`` (fp32) for a handful of small ranges, one launch per 16."""
    l = lib()
    l.dtg_pull_ranges.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                  ctypes.c_int, ctypes.c_void_p]
    l.dtg_pull_ranges.restype = ctypes.c_int
    ranges = [(int(a), int(b)) for a, b in ranges if b > a]
    for i in range(0, len(ranges), 16):
        part = ranges[i:i + 16]
        lo = (ctypes.c_longlong * len(part))(*[a for a, _ in part])
        hi = (ctypes.c_longlong * len(part))(*[b for _, b in part])
        _check(l.dtg_pull_ranges(src_f32.data_ptr(), dst_f32.data_ptr(), lo, hi, len(part), _stream()),
               "dtg_pull_ranges")


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


def softmax_xent_bf16_(logits: torch.Tensor, labels: torch.Tensor, loss_sum: torch.Tensor,
                       num_classes: int, scale: float) -> None:
    """In place on bf16 ``logits`` ``[B, ld]``: ``loss_sum += scale * sum(-log softmax[label])`` and
    ``logits <- scale * (softmax - onehot)``; columns ``>= num_classes`` are padding."""
    _need(logits, torch.bfloat16, "logits")
    _need(labels, torch.int64, "labels")
    l = lib()
    l.dtg_softmax_xent_bf16.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                        ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_void_p]
    l.dtg_softmax_xent_bf16.restype = ctypes.c_int
    _check(l.dtg_softmax_xent_bf16(logits.data_ptr(), labels.data_ptr(), loss_sum.data_ptr(),
                                   logits.shape[0], int(num_classes), logits.stride(0), float(scale),
                                   _stream()), "dtg_softmax_xent_bf16")


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


# --------------------------------------------------------------------------- #
# batched GEMM + transformer kernels
# --------------------------------------------------------------------------- #
def gemm_batched(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, *, a_mn: bool = False,
                 b_mn: bool = False, alpha: float = 1.0, mode: Optional[int] = None,
                 force_bn: int = 0) -> torch.Tensor:
    """``out[b2, b1] (op)= alpha * A[b2, b1] . B[b2, b1]`` for 4-D strided views.

    ``a``: ``[nb2, nb1, M, K]`` (or ``[nb2, nb1, K, M]`` with ``a_mn``); ``b``: ``[nb2, nb1, N,
    K]`` (or ``[.., K, N]`` with ``b_mn``); ``out``: ``[nb2, nb1, M, N]``.  Inner strides must
    be 1; everything else may be an arbitrary (16-byte multiple) stride, so per-head views
    of a packed QKV buffer work without copies.
    """
    _need(a, torch.bfloat16, "a")
    _need(b, torch.bfloat16, "b")
    if a.dim() != 4 or b.dim() != 4 or out.dim() != 4:
        raise ValueError("gemm_batched takes 4-D views [nb2, nb1, rows, inner]")
    if a.stride(3) != 1 or b.stride(3) != 1 or out.stride(3) != 1:
        raise ValueError("inner strides must be 1")
    nb2, nb1 = a.shape[0], a.shape[1]
    if a_mn:
        K, M = a.shape[2], a.shape[3]
    else:
        M, K = a.shape[2], a.shape[3]
    if b_mn:
        Kb, N = b.shape[2], b.shape[3]
    else:
        N, Kb = b.shape[2], b.shape[3]
    if K != Kb or tuple(b.shape[:2]) != (nb2, nb1) or tuple(out.shape) != (nb2, nb1, M, N):
        raise ValueError("batched gemm shape mismatch: a %s b %s out %s" % (
            tuple(a.shape), tuple(b.shape), tuple(out.shape)))
    if mode is None:
        mode = EPI_BF16 if out.dtype == torch.bfloat16 else EPI_F32
    _need(out, torch.bfloat16 if mode == EPI_BF16 else torch.float32, "out")
    rc = lib().dtg_gemm_bf16_batched(
        a.data_ptr(), b.data_ptr(), M, N, K, a.stride(2), b.stride(2), int(a_mn), int(b_mn),
        out.data_ptr(), out.stride(2), float(alpha), int(mode), nb1, nb2, a.stride(1), a.stride(0),
        b.stride(1), b.stride(0), out.stride(1), out.stride(0), int(force_bn), _stream())
    _check(rc, "dtg_gemm_bf16_batched")
    return out


def layernorm_fwd(x, gamma, beta, y, mean, rstd, eps: float = 1e-5) -> None:
    rows, cols = x.shape
    _check(lib().dtg_layernorm_fwd(x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), y.data_ptr(),
                                   _ptr(mean), _ptr(rstd), rows, cols, float(eps), _stream()),
           "dtg_layernorm_fwd")


def layernorm_bwd(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, alpha: float = 1.0,
                  dx_add: Optional[torch.Tensor] = None) -> None:
    """``dx = LN'(dy) (+ dx_add)``; ``dgamma/dbeta += alpha * (...)`` (reduction adds)."""
    rows, cols = x.shape
    _check(lib().dtg_layernorm_bwd(dy.data_ptr(), x.data_ptr(), gamma.data_ptr(), mean.data_ptr(),
                                   rstd.data_ptr(), _ptr(dx_add), dx.data_ptr(), _ptr(dgamma),
                                   _ptr(dbeta), rows, cols, float(alpha), _stream()),
           "dtg_layernorm_bwd")


def embedding_fwd(ids, wte, wpe, out, seq_len: int) -> None:
    rows, dim = out.shape
    _check(lib().dtg_embedding_fwd(ids.data_ptr(), wte.data_ptr(), _ptr(wpe), out.data_ptr(), rows,
                                   seq_len, dim, _stream()), "dtg_embedding_fwd")


def embedding_bwd(ids, dout, dwte, dwpe, seq_len: int, alpha: float = 1.0) -> None:
    rows, dim = dout.shape
    _check(lib().dtg_embedding_bwd(ids.data_ptr(), dout.data_ptr(), _ptr(dwte), _ptr(dwpe), rows,
                                   seq_len, dim, float(alpha), _stream()), "dtg_embedding_bwd")


def attn_softmax_fwd(S, P, T: int, scale: float, causal: bool) -> None:
    _check(lib().dtg_attn_softmax_fwd(S.data_ptr(), P.data_ptr(), S.numel() // T, T, float(scale),
                                      int(causal), _stream()), "dtg_attn_softmax_fwd")


def attn_softmax_bwd(P, dP, dS, T: int, scale: float) -> None:
    _check(lib().dtg_attn_softmax_bwd(P.data_ptr(), dP.data_ptr(), dS.data_ptr(), P.numel() // T,
                                      T, float(scale), _stream()), "dtg_attn_softmax_bwd")


def add_bf16(a, b, out) -> None:
    _check(lib().dtg_add_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.numel(), _stream()),
           "dtg_add_bf16")

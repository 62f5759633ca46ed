"""Attention on the batched tcgen05 GEMM with fused softmax epilogues.

Scores never reach HBM in fp32.  Forward: ``QK^T`` is computed twice -- the first pass'
epilogue keeps only per-(row, 128-key tile) running max / sum-exp statistics, a tiny merge
kernel turns them into ``m[row]`` and ``1/l[row]``, the second pass' epilogue writes the
normalised probabilities ``P`` straight in bf16 (the extra GEMM is ~1 % of a layer's FLOPs
and replaces ~1 GB of fp32 score traffic per layer).  Backward: ``dS = scale * P * (dO V^T -
D)`` is the epilogue of the ``dO V^T`` GEMM with ``D = rowsum(dO * O)``.  With ``causal`` the
tiles above the diagonal are skipped by every warp role; ``P`` / ``dS`` buffers must be
zero-initialised once because those tiles are never written.
"""
from __future__ import annotations

import ctypes

from . import _check, _stream, lib

EPI_ATTN_STATS, EPI_ATTN_P, EPI_ATTN_DS = 5, 6, 7
LOG2E = 1.4426950408889634
_declared = False


def _l():
    global _declared
    l = lib()
    if not _declared:
        vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float
        l.dtg_gemm_bf16_batched_ex.argtypes = [vp, vp, i32, i32, i32, i64, i64, i32, i32, vp, i64,
                                               f32, i32, i32, i32, i64, i64, i64, i64, i64, i64,
                                               vp, vp, vp, vp, f32, i32, vp]
        l.dtg_attn_stats_merge.argtypes = [vp, vp, vp, i64, i32, i32, i32, vp]
        l.dtg_attn_rowdot.argtypes = [vp, vp, vp, i32, i32, i32, i32, vp]
        for n in ("dtg_gemm_bf16_batched_ex", "dtg_attn_stats_merge", "dtg_attn_rowdot"):
            getattr(l, n).restype = i32
        _declared = True
    return l


def _call(a, b, out_ptr, out_strides, M, N, K, mode, alpha, rowv1, rowv2, stats, pmat, scale2, causal):
    nb2, nb1 = a.shape[0], a.shape[1]
    os2, os1, ldo = out_strides
    rc = _l().dtg_gemm_bf16_batched_ex(
        a.data_ptr(), b.data_ptr(), M, N, K, a.stride(2), b.stride(2), 0, 0, out_ptr, ldo,
        float(alpha), mode, nb1, nb2, a.stride(1), a.stride(0), b.stride(1), b.stride(0), os1, os2,
        None if rowv1 is None else rowv1.data_ptr(), None if rowv2 is None else rowv2.data_ptr(),
        None if stats is None else stats.data_ptr(), None if pmat is None else pmat.data_ptr(),
        float(scale2), int(causal), _stream())
    _check(rc, "dtg_gemm_bf16_batched_ex")


def softmax_probs(q, k, P, stats, rowm, rowinvl, scale: float, causal: bool) -> None:
    """``P[b,h] = softmax(scale * q[b,h] k[b,h]^T)`` (bf16, [B,H,T,T] contiguous, zero-initialised).
    ``q``/``k``: [B, H, T, hd] strided views; ``stats``: fp32 [B*H*T, ceil(T/128), 2];
    ``rowm``/``rowinvl``: fp32 [B*H*T]."""
    B, H, T, hd = q.shape
    s2 = scale * LOG2E
    tiles_n = (T + 127) // 128
    _call(q, k, None, (0, 0, T), T, T, hd, EPI_ATTN_STATS, 1.0, None, None, stats, None, s2, causal)
    _check(_l().dtg_attn_stats_merge(stats.data_ptr(), rowm.data_ptr(), rowinvl.data_ptr(),
                                     B * H * T, tiles_n, T, int(causal), _stream()),
           "dtg_attn_stats_merge")
    _call(q, k, P.data_ptr(), (P.stride(0), P.stride(1), P.stride(2)), T, T, hd, EPI_ATTN_P, 1.0,
          rowm, rowinvl, None, None, s2, causal)


def rowdot(dO, O, D, B, T, H, hd) -> None:
    """``D[b,h,t] = sum_c dO[b*T+t, h*hd+c] * O[...]`` (fp32 [B*H*T])."""
    _check(_l().dtg_attn_rowdot(dO.data_ptr(), O.data_ptr(), D.data_ptr(), B, T, H, hd, _stream()),
           "dtg_attn_rowdot")


def dscores(dctx, v, P, D, dS, scale: float, causal: bool) -> None:
    """``dS[b,h] = scale * P * (dctx[b,h] v[b,h]^T - D)`` (bf16, zero-initialised buffer)."""
    B, H, T, hd = dctx.shape
    _call(dctx, v, dS.data_ptr(), (dS.stride(0), dS.stride(1), dS.stride(2)), T, T, hd,
          EPI_ATTN_DS, scale, D, None, None, P, 0.0, causal)


_fwd_declared = False


def fused_forward(q, k, v, ctx, rowm, rowinvl, scale: float, causal: bool, P=None) -> None:
    """One kernel (``csrc/attn_sm100.cu``): ``ctx[b*T+t, h*64:(h+1)*64] = softmax(scale q k^T) v``.

    ``q``/``k``/``v``: [B, H, T, 64] strided views of one packed buffer (same strides);
    ``ctx``: [B*T, H*64] bf16; ``rowm``/``rowinvl``: fp32 [B*H*T] (saved for the backward pass);
    ``P``: optional zero-initialised bf16 [B, H, T, T] that receives the probabilities by TMA store.
    """
    global _fwd_declared
    l = _l()
    if not _fwd_declared:
        vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float
        l.dtg_attn_fwd.argtypes = [vp, vp, vp, i64, i64, i64, i32, i32, i32, i32, vp, vp, i64, vp, vp,
                                   f32, i32, vp]
        l.dtg_attn_fwd.restype = i32
        _fwd_declared = True
    B, H, T, hd = q.shape
    if k.stride() != q.stride() or v.stride() != q.stride() or q.stride(3) != 1:
        raise ValueError("q, k, v must be views with identical strides and unit inner stride")
    _check(l.dtg_attn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), q.stride(2), q.stride(1),
                          q.stride(0), B, H, T, hd, None if P is None else P.data_ptr(),
                          ctx.data_ptr(), ctx.stride(0), rowm.data_ptr(), rowinvl.data_ptr(),
                          float(scale), int(causal), _stream()), "dtg_attn_fwd")


_bwd_declared = False


def fused_backward(q, k, v, dk, dv, dO, dS, rowm, rowinvl, D, scale: float, causal: bool) -> None:
    """``dk``, ``dv`` and ``dS`` in one kernel; P is recomputed from (``rowm``, ``rowinvl``).

    ``q``/``k``/``v`` and ``dk``/``dv``: [B, H, T, 64] views with identical strides; ``dO``:
    [B, H, T, 64] view; ``dS``: bf16 [B, H, T, T] (zero-initialised when ``causal``); ``D`` =
    ``rowdot(dO, O)``.  ``dq = dS k`` is a batched GEMM on the stored ``dS``.
    """
    global _bwd_declared
    l = _l()
    if not _bwd_declared:
        vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float
        l.dtg_attn_bwd.argtypes = [vp, vp, vp, vp, vp, i64, i64, i64, vp, i64, i64, i64, i32, i32, i32,
                                   i32, vp, vp, vp, vp, f32, i32, vp]
        l.dtg_attn_bwd.restype = i32
        _bwd_declared = True
    B, H, T, hd = q.shape
    for t in (k, v, dk, dv):
        if t.stride() != q.stride():
            raise ValueError("q, k, v, dk, dv must be views with identical strides")
    _check(l.dtg_attn_bwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), dk.data_ptr(), dv.data_ptr(),
                          q.stride(2), q.stride(1), q.stride(0), dO.data_ptr(), dO.stride(2),
                          dO.stride(1), dO.stride(0), B, H, T, hd, dS.data_ptr(), rowm.data_ptr(),
                          rowinvl.data_ptr(), D.data_ptr(), float(scale), int(causal), _stream()),
           "dtg_attn_bwd")

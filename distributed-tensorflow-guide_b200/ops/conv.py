"""ctypes wrappers for ``ops/csrc/conv_kernels.cu`` (NHWC bf16): im2col / col2im,
BatchNorm with batch statistics (+ReLU, +residual), max and global-average pooling."""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _check, _ptr, _stream, lib

_declared = False


def _l():
    global _declared
    l = lib()
    if not _declared:
        vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float
        l.dtg_im2col.argtypes = [vp, vp] + [i32] * 8 + [vp]
        l.dtg_col2im.argtypes = [vp, vp, vp] + [i32] * 8 + [vp]
        l.dtg_bn_stats.argtypes = [vp, vp, vp, i64, i32, vp]
        l.dtg_bn_apply.argtypes = [vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, i32, vp]
        l.dtg_bn_bwd.argtypes = [vp] * 12 + [i64, i32, f32, i32, f32, vp]
        l.dtg_maxpool_fwd.argtypes = [vp, vp] + [i32] * 7 + [vp]
        l.dtg_maxpool_bwd.argtypes = [vp, vp, vp, vp] + [i32] * 7 + [vp]
        l.dtg_avgpool_fwd.argtypes = [vp, vp, i32, i32, i32, vp]
        l.dtg_avgpool_bwd.argtypes = [vp, vp, i32, i32, i32, vp]
        l.dtg_conv_gemm_bf16.argtypes = [vp] + [i32] * 8 + [vp, i64, i32, vp, i64, vp, f32, i32, i32, vp]
        l.dtg_conv_gemm_bf16.restype = i32
        for n in ("dtg_im2col", "dtg_col2im", "dtg_bn_stats", "dtg_bn_apply", "dtg_bn_bwd",
                  "dtg_maxpool_fwd", "dtg_maxpool_bwd", "dtg_avgpool_fwd", "dtg_avgpool_bwd"):
            getattr(l, n).restype = i32
        _declared = True
    return l


def out_size(n: int, k: int, stride: int, pad: int) -> int:
    return (n + 2 * pad - k) // stride + 1


def implicit_ok(C: int, k: int) -> bool:
    """Can this convolution run as an implicit GEMM (TMA im2col loads)?  64-channel k-blocks."""
    return C % 64 == 0 and k >= 1


def conv_gemm_fwd(x: torch.Tensor, w: torch.Tensor, out: torch.Tensor, N, H, W, C, k, stride, pad,
                  bias: Optional[torch.Tensor] = None, force_bn: int = 0) -> None:
    """``out[N*Ho*Wo, Cout] = im2col(x) . w[Cout, k*k*C]^T`` without materialising the patch matrix:
    the A operand of the tcgen05 GEMM is gathered from ``x`` (NHWC bf16) by TMA im2col loads."""
    from . import EPI_BF16, EPI_F32
    mode = EPI_BF16 if out.dtype == torch.bfloat16 else EPI_F32
    _check(_l().dtg_conv_gemm_bf16(x.data_ptr(), N, H, W, C, k, stride, pad, 1, w.data_ptr(),
                                   w.stride(0), w.shape[0], out.data_ptr(), out.stride(0), _ptr(bias),
                                   1.0, mode, force_bn, _stream()), "dtg_conv_gemm_bf16")


def conv_gemm_wgrad(dy: torch.Tensor, x: torch.Tensor, target: torch.Tensor, N, H, W, C, k, stride,
                    pad, alpha: float = 1.0, mode: Optional[int] = None, force_bn: int = 0) -> None:
    """``target[Cout, k*k*C] (+)= alpha * dy[N*Ho*Wo, Cout]^T . im2col(x)``: the B operand is gathered
    by TMA im2col loads; ``mode`` is any fp32 sink of ``ops.gemm`` (store, +=, red.add push)."""
    from . import EPI_F32
    _check(_l().dtg_conv_gemm_bf16(x.data_ptr(), N, H, W, C, k, stride, pad, 2, dy.data_ptr(),
                                   dy.stride(0), dy.shape[1], target.data_ptr(), target.stride(0),
                                   None, float(alpha), EPI_F32 if mode is None else int(mode),
                                   force_bn, _stream()), "dtg_conv_gemm_bf16")


def conv_gemm_dgrad(dy: torch.Tensor, w: torch.Tensor, dx: torch.Tensor, N, H, W, C, k, pad) -> None:
    """Data gradient of a STRIDE-1 k x k convolution as an implicit GEMM over ``dy``:
    ``dx[N*H*W, C] = im2col_{pad' = k-1-pad}(dy[N, Ho, Wo, Cout]) . Wf[C, k*k*Cout]^T`` with the
    filter flipped and its channel axes swapped (``w`` is ``[Cout, k*k*C]``).  No ``dcol`` matrix,
    no ``col2im`` pass."""
    Cout = w.shape[0]
    Ho, Wo = out_size(H, k, 1, pad), out_size(W, k, 1, pad)
    wf = w.view(Cout, k, k, C).flip(1, 2).permute(3, 1, 2, 0).reshape(C, k * k * Cout).contiguous()
    conv_gemm_fwd(dy, wf, dx, N, Ho, Wo, Cout, k, 1, k - 1 - pad)


def im2col(x: torch.Tensor, col: torch.Tensor, N, H, W, C, kh, kw, stride, pad) -> None:
    _check(_l().dtg_im2col(x.data_ptr(), col.data_ptr(), N, H, W, C, kh, kw, stride, pad, _stream()),
           "dtg_im2col")


def col2im(dcol, dx, N, H, W, C, kh, kw, stride, pad, dx_add: Optional[torch.Tensor] = None) -> None:
    _check(_l().dtg_col2im(dcol.data_ptr(), _ptr(dx_add), dx.data_ptr(), N, H, W, C, kh, kw, stride,
                           pad, _stream()), "dtg_col2im")


def bn_stats(x, sums, sumsq, zero: bool = True) -> None:
    """``zero=False``: the caller has already cleared the accumulators (e.g. all layers' at once)."""
    rows, C = x.shape
    if zero:
        sums.zero_(); sumsq.zero_()
    _check(_l().dtg_bn_stats(x.data_ptr(), sums.data_ptr(), sumsq.data_ptr(), rows, C, _stream()),
           "dtg_bn_stats")


def bn_apply(x, sums, sumsq, gamma, beta, y, relu: bool, residual=None, eps: float = 1e-5) -> None:
    rows, C = x.shape
    _check(_l().dtg_bn_apply(x.data_ptr(), sums.data_ptr(), sumsq.data_ptr(), gamma.data_ptr(),
                             beta.data_ptr(), _ptr(residual), y.data_ptr(), rows, C, float(eps),
                             int(relu), _stream()), "dtg_bn_apply")


def bn_bwd(dy, x, y, sums, sumsq, gamma, scratch, dx, dgamma, dbeta, relu: bool, alpha: float = 1.0,
           dres=None, eps: float = 1e-5) -> None:
    """``scratch``: fp32 [2, C].  dgamma/dbeta receive ``alpha * grad`` as reduction adds."""
    rows, C = x.shape
    _check(_l().dtg_bn_bwd(dy.data_ptr(), x.data_ptr(), _ptr(y), sums.data_ptr(), sumsq.data_ptr(),
                           gamma.data_ptr(), scratch[0].data_ptr(), scratch[1].data_ptr(),
                           dx.data_ptr(), _ptr(dres), _ptr(dgamma), _ptr(dbeta), rows, C, float(eps),
                           int(relu), float(alpha), _stream()), "dtg_bn_bwd")


def maxpool_fwd(x, y, N, H, W, C, k, stride, pad) -> None:
    _check(_l().dtg_maxpool_fwd(x.data_ptr(), y.data_ptr(), N, H, W, C, k, stride, pad, _stream()),
           "dtg_maxpool_fwd")


def maxpool_bwd(x, y, dy, dx, N, H, W, C, k, stride, pad) -> None:
    _check(_l().dtg_maxpool_bwd(x.data_ptr(), y.data_ptr(), dy.data_ptr(), dx.data_ptr(), N, H, W, C,
                                k, stride, pad, _stream()), "dtg_maxpool_bwd")


def avgpool_fwd(x, y, N, HW, C) -> None:
    _check(_l().dtg_avgpool_fwd(x.data_ptr(), y.data_ptr(), N, HW, C, _stream()), "dtg_avgpool_fwd")


def avgpool_bwd(dy, dx, N, HW, C) -> None:
    _check(_l().dtg_avgpool_bwd(dy.data_ptr(), dx.data_ptr(), N, HW, C, _stream()), "dtg_avgpool_bwd")

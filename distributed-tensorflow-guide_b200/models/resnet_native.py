"""sm_100a implementation of the ResNet family (BASELINE.json config 3).

NHWC bf16 activations.  Every convolution is a tcgen05 GEMM: 1x1/stride-1 directly on
the ``[N*H*W, Cin]`` activation matrix, everything else after an ``im2col`` gather;
weight gradients are the MN-major GEMM ``dY^T . col`` with the usual gradient sinks
(store / window accumulate / push into the ps shard), data gradients are ``dY . W``
followed by ``col2im``.  BatchNorm (batch statistics) + ReLU + the residual add are one
elementwise kernel after a column-statistics kernel; their backward is the mirrored
pair.  Plugs into ``GpuWorker`` like the other native models.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from .. import ops
from ..ops import conv as C
from ..parallel.flat import FlatLayout
from .mlp_native import GradSink
from .resnet import CIN_PAD, ConvSpec, ResNet, ResNetConfig

__all__ = ["NativeResNet"]


import os as _os
_IMPLICIT = _os.environ.get("DTG_CONV_IMPLICIT", "1") != "0"


def _pad8(n):
    return -(-n // 8) * 8


class _Unit:
    """One conv+BN(+ReLU)(+residual) with its saved tensors."""

    def __init__(self, cv: ConvSpec, N: int, H: int, W: int, dev):
        self.cv, self.N, self.H, self.W = cv, N, H, W
        self.Ho = C.out_size(H, cv.k, cv.stride, cv.pad)
        self.Wo = C.out_size(W, cv.k, cv.stride, cv.pad)
        self.rows_in, self.rows = N * H * W, N * self.Ho * self.Wo
        self.K = cv.k * cv.k * cv.cin
        self.direct = cv.k == 1 and cv.stride == 1
        bf = dict(device=dev, dtype=torch.bfloat16)
        self.raw = torch.empty(self.rows, cv.cout, **bf)      # conv output (pre-BN)
        self.out = torch.empty(self.rows, cv.cout, **bf)      # after BN (+res) (+ReLU)
        self.sums = torch.zeros(cv.cout, device=dev)
        self.sumsq = torch.zeros(cv.cout, device=dev)
        self.x: Optional[torch.Tensor] = None
        self.relu_out = False
        # The patch matrix of a k > 1 (or strided) convolution is kept from forward to backward:
        # 6.4 GB over ResNet-50's 3x3 layers at batch 256 -- nothing next to 180 GB of HBM, and it
        # saves the second im2col pass (the same bytes written again) in every backward.
        # ... unless the convolution runs as an implicit GEMM (64-channel k-blocks gathered by TMA
        # im2col loads): then no patch matrix exists at all, forward or backward.
        self.implicit = (not self.direct) and C.implicit_ok(cv.cin, cv.k) and _IMPLICIT
        self.col = None if (self.direct or self.implicit) else torch.empty(self.rows, self.K, **bf)


class NativeResNet(ResNet):
    def __init__(self, cfg: ResNetConfig, batch_size: int, device="cuda", num_shards: int = 1,
                 sharding: str = "round_robin"):
        super().__init__(cfg, compute_dtype=torch.bfloat16)
        self.device = torch.device(device)
        self.B = int(batch_size)
        self.classes_pad = _pad8(cfg.classes)
        specs = []
        for s in self.param_specs():
            stored = s.shape
            if s.name.endswith("logits/weights"):
                stored = (self.classes_pad, s.shape[1])
            elif s.name.endswith("logits/biases"):
                stored = (self.classes_pad,)
            specs.append((s.name, s.shape, stored))
        self.layout = FlatLayout(specs, num_shards, sharding)
        dev, n = self.device, self.layout.total
        self.p32 = torch.zeros(n, device=dev)
        self.p16 = torch.zeros(n, device=dev, dtype=torch.bfloat16)
        self.g32 = torch.zeros(n, device=dev)
        self.after_layer = None
        N, S = self.B, cfg.image
        bf = dict(device=dev, dtype=torch.bfloat16)
        # ---- build the unit graph with shapes
        self.stem = _Unit(self.stem_conv(), N, S, S, dev)
        hp = C.out_size(self.stem.Ho, 3, 2, 1)
        self.pool_out = torch.empty(N * hp * hp, cfg.stem, **bf)
        self.blocks: List[dict] = []
        h = hp
        max_col, max_act = self.stem.rows * self.stem.K, self.stem.rows * cfg.stem
        for scope, convs, sc in self.units():
            u1 = _Unit(convs[0], N, h, h, dev)
            u2 = _Unit(convs[1], N, u1.Ho, u1.Wo, dev)
            u3 = _Unit(convs[2], N, u2.Ho, u2.Wo, dev)
            usc = _Unit(sc, N, h, h, dev) if sc else None
            self.blocks.append({"scope": scope, "u": [u1, u2, u3], "sc": usc, "hin": h})
            for u in [u1, u2, u3] + ([usc] if usc else []):
                if not u.direct:
                    max_col = max(max_col, u.rows * u.K)
                max_act = max(max_act, u.rows * u.cv.cout, u.rows_in * u.cv.cin)
            h = u3.Ho
        self.h_last = h
        # every layer's BatchNorm accumulators in one buffer: cleared by ONE memset per forward
        units = [self.stem] + [u for b in self.blocks for u in b["u"] + ([b["sc"]] if b["sc"] else [])]
        self._bn_stat = torch.zeros(2 * sum(u.cv.cout for u in units), device=dev)
        o = 0
        for u in units:
            u.sums = self._bn_stat[o:o + u.cv.cout]
            u.sumsq = self._bn_stat[o + u.cv.cout:o + 2 * u.cv.cout]
            o += 2 * u.cv.cout
        feat = 4 * cfg.widths[-1]
        self.col = torch.empty(max_col, **bf)                  # im2col / dcol scratch
        self.pooled = torch.empty(N, feat, **bf)
        self.logits = torch.zeros(N, self.classes_pad, device=dev)
        self.dlogits = torch.zeros(N, self.classes_pad, **bf)
        self.dpooled = torch.empty(N, feat, **bf)
        self.gbuf = [torch.empty(max_act, **bf) for _ in range(4)]   # gradient ping-pong
        self.bn_scratch = torch.zeros(2, 4 * cfg.widths[-1], device=dev)
        self.x_in = torch.empty(N * S * S, CIN_PAD, **bf)
        self.loss_buf = torch.zeros(1, device=dev)
        self._bind_views()

    # ------------------------------------------------------------- plumbing
    def _views(self, flat):
        return {n: self.layout.view(flat, n) for n in self.layout.order}

    def _bind_views(self):
        self.w16, self.w32 = self._views(self.p16), self._views(self.p32)
        self.local_sink = GradSink("store", self._views(self.g32))

    def rebind(self, p32, p16, g32):
        p32.copy_(self.p32); p16.copy_(self.p16); g32.zero_()
        self.p32, self.p16, self.g32 = p32, p16, g32
        self._bind_views()

    def sink(self, mode, flat_or_views, alpha: float = 1.0, remote=()) -> GradSink:
        views = flat_or_views if isinstance(flat_or_views, dict) else self._views(flat_or_views)
        return GradSink(mode, views, alpha, remote)

    def load_params(self, values):
        self.p32.zero_()
        self.layout.pack(self.p32, {k: v.to(self.device) for k, v in values.items()})
        ops.cast_f32_bf16(self.p32, self.p16)

    def refresh_bf16(self):
        ops.cast_f32_bf16(self.p32, self.p16)

    def init_on_device(self, seed: int = 0) -> int:
        """Draw every parameter on the GPU, straight into the flat fp32 buffer (``ops/init.py``:
        counter-based, so equal seeds give bit-identical replicas on every rank)."""
        from ..ops import init as dinit
        n = dinit.device_init(self, self.layout, self.p32, seed)
        self.refresh_bf16()
        return n

    def export_params(self):
        return {k: v.clone() for k, v in self.layout.unpack(self.p32).items()}

    def export_grads(self):
        return {k: v.clone() for k, v in self.layout.unpack(self.g32).items()}

    def synthetic_batch(self, batch_size=None, device=None, generator=None):
        return super().synthetic_batch(batch_size or self.B, self.device, generator)

    # --------------------------------------------------------------- forward
    def _conv_fwd(self, u: _Unit, x: torch.Tensor, residual=None, relu=None):
        cv = u.cv
        u.x = x
        if u.implicit:
            C.conv_gemm_fwd(x, self.w16[cv.name + "/weights"], u.raw, u.N, u.H, u.W, cv.cin, cv.k,
                            cv.stride, cv.pad)
        else:
            if u.direct:
                a = x.view(u.rows, cv.cin)
            else:
                a = u.col
                C.im2col(x, a, u.N, u.H, u.W, cv.cin, cv.k, cv.k, cv.stride, cv.pad)
            ops.gemm(a, self.w16[cv.name + "/weights"], u.raw)
        C.bn_stats(u.raw, u.sums, u.sumsq, zero=False)
        u.relu_out = cv.relu if relu is None else relu
        C.bn_apply(u.raw, u.sums, u.sumsq, self.w32[cv.name + "/BatchNorm/gamma"],
                   self.w32[cv.name + "/BatchNorm/beta"], u.out, u.relu_out, residual)
        return u.out

    def forward(self, x: torch.Tensor, labels: torch.Tensor) -> None:
        cfg, N = self.cfg, self.B
        self._bn_stat.zero_()
        if x.shape[-1] == 3:
            # 3-channel images: pad to 8 channels (zeros) so rows are 16-byte multiples
            xin = self.x_in.view(N, cfg.image, cfg.image, CIN_PAD)
            xin[..., 3:].zero_()
            xin[..., :3].copy_(x)
        elif x.dtype == torch.bfloat16:
            self.x_in.view(-1).copy_(x.reshape(-1))
        else:
            ops.cast_f32_bf16(x.contiguous().view(-1), self.x_in.view(-1))
        h = self._conv_fwd(self.stem, self.x_in)
        C.maxpool_fwd(h, self.pool_out, N, self.stem.Ho, self.stem.Wo, cfg.stem, 3, 2, 1)
        h = self.pool_out
        for blk in self.blocks:
            u1, u2, u3 = blk["u"]
            idn = h if blk["sc"] is None else self._conv_fwd(blk["sc"], h)
            a = self._conv_fwd(u1, h)
            b = self._conv_fwd(u2, a)
            h = self._conv_fwd(u3, b, residual=idn, relu=True)
        feat = 4 * cfg.widths[-1]
        C.avgpool_fwd(h, self.pooled, N, self.h_last * self.h_last, feat)
        ops.gemm(self.pooled, self.w16[cfg.name + "/logits/weights"], self.logits,
                 bias=self.w32[cfg.name + "/logits/biases"], mode=ops.EPI_F32)
        self.loss_buf.zero_()
        ops.softmax_xent(self.logits, labels, self.dlogits, self.loss_buf, cfg.classes, 1.0 / N)

    # -------------------------------------------------------------- backward
    def _conv_bwd(self, u: _Unit, dy: torch.Tensor, sink: GradSink, dx: Optional[torch.Tensor],
                  dres: Optional[torch.Tensor] = None, dx_add: Optional[torch.Tensor] = None):
        """``dy``: grad of ``u.out``.  Writes grad of the conv input into ``dx`` (None for the
        stem), the masked upstream grad into ``dres`` (residual branch), parameter grads into
        the sink.  ``dx_add`` is summed into ``dx``."""
        cv = u.cv
        g_raw = self.gbuf[3][:u.rows * cv.cout].view(u.rows, cv.cout)
        C.bn_bwd(dy, u.raw, u.out, u.sums, u.sumsq, self.w32[cv.name + "/BatchNorm/gamma"],
                 self._scratch(cv.cout),
                 g_raw, sink.targets[cv.name + "/BatchNorm/gamma"],
                 sink.targets[cv.name + "/BatchNorm/beta"], u.relu_out, sink.alpha, dres)
        wname = cv.name + "/weights"
        if u.implicit:
            C.conv_gemm_wgrad(g_raw, u.x, sink.targets[wname], u.N, u.H, u.W, cv.cin, cv.k, cv.stride,
                              cv.pad, alpha=sink.alpha, mode=sink.epi_for(wname))
        else:
            if u.direct:
                a = u.x.view(u.rows, cv.cin)
            else:
                a = u.col                                  # written by this step's forward
            ops.gemm(g_raw, a, sink.targets[wname], a_mn=True, b_mn=True, alpha=sink.alpha,
                     mode=sink.epi_for(wname))
        if dx is None:
            return
        if u.direct:
            ops.gemm(g_raw, self.w16[wname], dx.view(u.rows, cv.cin), b_mn=True,
                     residual=None if dx_add is None else dx_add.view(u.rows, cv.cin))
        elif (u.implicit and cv.stride == 1 and cv.k > 1 and dx_add is None
              and C.implicit_ok(cv.cout, cv.k)):
            # stride-1 k x k: the data gradient is itself a convolution of dY (flipped filter)
            C.conv_gemm_dgrad(g_raw, self.w16[wname], dx.view(u.rows_in, cv.cin), u.N, u.H, u.W,
                              cv.cin, cv.k, cv.pad)
        else:
            dcol = self.col[:u.rows * u.K].view(u.rows, u.K)
            ops.gemm(g_raw, self.w16[wname], dcol, b_mn=True)
            C.col2im(dcol, dx, u.N, u.H, u.W, cv.cin, cv.k, cv.k, cv.stride, cv.pad, dx_add)

    def _scratch(self, c):
        return [self.bn_scratch[0, :c], self.bn_scratch[1, :c]]

    def backward(self, sink: GradSink) -> None:
        cfg, N = self.cfg, self.B
        if sink.mode == "store":
            sink.zero_reduction_targets()
        feat = 4 * cfg.widths[-1]
        wl, bl = cfg.name + "/logits/weights", cfg.name + "/logits/biases"
        ops.gemm(self.dlogits, self.pooled, sink.targets[wl], a_mn=True, b_mn=True,
                 alpha=sink.alpha, mode=sink.epi_for(wl))
        ops.bias_grad(self.dlogits, sink.targets[bl], sink.alpha)
        ops.gemm(self.dlogits, self.w16[wl], self.dpooled, b_mn=True)
        ga, gb, gc = self.gbuf[0], self.gbuf[1], self.gbuf[2]
        last = self.blocks[-1]["u"][2]
        dh = ga[:last.rows * last.cv.cout].view(last.rows, last.cv.cout)
        C.avgpool_bwd(self.dpooled, dh, N, self.h_last * self.h_last, feat)
        cur, f1, f2 = ga, gb, gc
        for li in range(len(self.blocks) - 1, -1, -1):
            blk = self.blocks[li]
            u1, u2, u3 = blk["u"]
            sc = blk["sc"]
            cin, hin = u1.cv.cin, blk["hin"]
            rows_in = N * hin * hin
            dh = cur[:u3.rows * u3.cv.cout].view(u3.rows, u3.cv.cout)
            # conv3 (+ residual join): dres = masked upstream grad for the shortcut branch
            dres = f1[:u3.rows * u3.cv.cout].view(u3.rows, u3.cv.cout)
            db = f2[:u3.rows_in * u3.cv.cin].view(u3.rows_in, u3.cv.cin)
            self._conv_bwd(u3, dh, sink, db, dres=dres)
            da = cur[:u2.rows_in * u2.cv.cin].view(u2.rows_in, u2.cv.cin)     # cur is free now
            self._conv_bwd(u2, db, sink, da)
            dx_main = f2[:rows_in * cin].view(rows_in, cin)
            if sc is None:
                # identity shortcut: dx = dx_main + dres (fused into the 1x1 dgrad epilogue)
                self._conv_bwd(u1, da, sink, dx_main, dx_add=dres)
                new_cur = f2
            else:
                self._conv_bwd(u1, da, sink, dx_main)
                dx = cur[:rows_in * cin].view(rows_in, cin)
                self._conv_bwd(sc, dres, sink, dx, dx_add=dx_main)
                new_cur = cur
            if new_cur is f2:
                cur, f1, f2 = f2, f1, cur
            if self.after_layer is not None:
                self.after_layer(li)
        # stem: maxpool backward then conv1 (no input gradient needed)
        dpool = cur[:self.pool_out.numel()].view_as(self.pool_out)
        dstem = f1[:self.stem.rows * cfg.stem].view(self.stem.rows, cfg.stem)
        C.maxpool_bwd(self.stem.out, self.pool_out, dpool, dstem, N, self.stem.Ho, self.stem.Wo,
                      cfg.stem, 3, 2, 1)
        self._conv_bwd(self.stem, dstem, sink, None)

    def layer_bounds(self):
        out = []
        for blk in self.blocks:
            names = []
            for u in blk["u"] + ([blk["sc"]] if blk["sc"] else []):
                names += [u.cv.name + s for s in ("/weights", "/BatchNorm/gamma", "/BatchNorm/beta")]
            es = [self.layout.entries[n] for n in names]
            out.append((min(e.offset for e in es),
                        max(e.offset + -(-e.numel // 64) * 64 for e in es)))
        return out

    def forward_backward_native(self, x, labels, sink: Optional[GradSink] = None) -> torch.Tensor:
        self.forward(x, labels)
        self.backward(sink or self.local_sink)
        return self.loss_buf

    def forward_backward(self, params, batch):
        self.load_params(params)
        x, y = batch
        loss = self.forward_backward_native(x, y).clone()
        return loss[0], self.export_grads()

    def local_apply(self, lr: float) -> None:
        ops.local_step(self.p32, self.g32, lr, p_bf16=self.p16)

    def train_step_local(self, x, y, lr: float) -> float:
        loss = self.forward_backward_native(x, y)
        self.local_apply(lr)
        return float(loss.item())

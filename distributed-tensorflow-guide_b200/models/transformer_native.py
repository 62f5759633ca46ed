"""sm_100a implementation of the transformer family (GPT-2 medium / BERT-large).

Every dense matmul -- QKV, output projection, MLP, LM head and their dgrad / wgrad -- runs on
the tcgen05 GEMM of ``ops``; bias, GELU, GELU', residual adds and the gradient sink are GEMM
epilogues (wgrads split over K when the output is small).  Attention with head_dim 64 is the
pair of fused kernels of ``ops/csrc/attn_sm100.cu``: scores and probabilities never reach HBM,
only the row statistics (max, 1/sum-exp) are kept per layer and ``P`` is recomputed in the
backward pass; ``dQ = dS K`` is one batched GEMM on the ``dS`` the backward kernel stores.
Other head sizes fall back to the batched score GEMM with softmax epilogues
(``ops/attn.py``) or, with ``fused_attn=False``, to the fully unfused path.  LayerNorm and the
embedding are the row-wise kernels of ``ops/csrc/transformer_kernels.cu``.

Both block orderings of ``models/transformer.py`` run on the same kernels: pre-LN (GPT-2) folds
the residual adds into the projection GEMMs' epilogues and the LayerNorm backward's ``dx_add``;
post-LN (BERT, ``cfg.post_ln``) normalises the epilogue's ``x + proj`` sum and carries the residual
gradient through the dgrad GEMMs' residual epilogue instead.

The model plugs into ``GpuWorker`` exactly like ``NativeMLP``: flat fp32/bf16 parameter
buffers, a gradient sink (store / accumulate / push into the ps shard) and a per-layer
callback that lets the synchronous schemes launch the fused reduce-scatter/update/
all-gather kernel for a layer's bucket while earlier layers are still in backward.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch

from .. import ops
from ..ops import attn
from ..parallel.flat import FlatLayout
from .mlp_native import GradSink
from .transformer import Transformer, TransformerConfig

__all__ = ["NativeTransformer"]


class NativeTransformer(Transformer):
    def __init__(self, cfg: TransformerConfig, batch_size: int, device="cuda", num_shards: int = 1,
                 sharding: str = "round_robin", fused_attn: bool = True):
        super().__init__(cfg, compute_dtype=torch.bfloat16)
        c = cfg
        if c.head_dim % 8 or c.dim % 8 or c.ff % 8:
            raise ValueError("dims must be multiples of 8 (TMA row pitch)")
        self.device = torch.device(device)
        self.B, self.T = int(batch_size), c.seq
        self.R = self.B * self.T
        specs = []
        for s in self.param_specs():
            stored = (c.vocab_padded, c.dim) if s.name == "model/wte" else s.shape
            specs.append((s.name, s.shape, stored))
        self.layout = FlatLayout(specs, num_shards, sharding)
        dev, n = self.device, self.layout.total
        self.p32 = torch.zeros(n, device=dev)
        self.p16 = torch.zeros(n, device=dev, dtype=torch.bfloat16)
        self.g32 = torch.zeros(n, device=dev)
        self.after_layer = None
        R, d, H, hd, T, B = self.R, c.dim, c.heads, c.head_dim, self.T, self.B
        bf = dict(device=dev, dtype=torch.bfloat16)
        f32 = dict(device=dev, dtype=torch.float32)
        L = c.layers
        # saved activations, per layer
        self.x = [torch.empty(R, d, **bf) for _ in range(L + 1)]     # block inputs (+ final)
        self.h1 = [torch.empty(R, d, **bf) for _ in range(L)]
        self.qkv = [torch.empty(R, 3 * d, **bf) for _ in range(L)]
        self.fused_attn = fused_attn
        self.flash = fused_attn and hd == 64     # tcgen05 attention kernels, P never stored
        # zero-initialised: causal tiles above the diagonal are skipped, never written
        self.P = None if self.flash else [torch.zeros(B, H, T, T, **bf) for _ in range(L)]
        self.ctx = [torch.empty(R, d, **bf) for _ in range(L)]
        self.x2 = [torch.empty(R, d, **bf) for _ in range(L)]
        self.h2 = [torch.empty(R, d, **bf) for _ in range(L)]
        self.z = [torch.empty(R, c.ff, **bf) for _ in range(L)]
        self.f = [torch.empty(R, c.ff, **bf) for _ in range(L)]
        self.stats = [[torch.empty(R, **f32) for _ in range(4)] for _ in range(L)]
        # pre-LN: output of the final LayerNorm; post-LN: the raw embedding sum (LN_emb's input)
        self.hf = torch.empty(R, d, **bf)
        self.stat_f = [torch.empty(R, **f32) for _ in range(2)]
        self.post_ln = bool(c.post_ln)
        # scratch shared by all layers
        if fused_attn:
            # softmax lives in the attention kernels / score-GEMM epilogues: only O(rows) fp32
            # statistics (row max and 1/sum-exp, kept per layer for the backward pass)
            self.S = None
            self.att_stats = None if self.flash else torch.empty(B * H * T, (T + 127) // 128, 2, **f32)
            self.att_m = [torch.empty(B * H * T, **f32) for _ in range(L if self.flash else 1)]
            self.att_il = [torch.empty(B * H * T, **f32) for _ in range(L if self.flash else 1)]
            self.att_D = torch.empty(B * H * T, **f32)
        else:
            self.S = torch.empty(B, H, T, T, **f32)      # scores / dP
        self.dS = torch.zeros(B, H, T, T, **bf)
        # bf16 logits; the loss kernel rewrites them in place as dlogits (one buffer, one pass)
        self.logits = torch.empty(R, c.vocab_padded, **bf)
        self.dlogits = self.logits
        self.dqkv = torch.empty(R, 3 * d, **bf)
        self.dff = torch.empty(R, c.ff, **bf)
        self.da = [torch.empty(R, d, **bf) for _ in range(3)]
        self.loss_buf = torch.zeros(1, **f32)
        self._bind_views()

    # ------------------------------------------------------------------ views
    def _views(self, flat):
        return {n: self.layout.view(flat, n) for n in self.layout.order}

    def _bind_views(self):
        self.w16 = self._views(self.p16)
        self.w32 = self._views(self.p32)
        self.local_sink = GradSink("store", self._views(self.g32))

    def rebind(self, p32, p16, g32):
        p32.copy_(self.p32); p16.copy_(self.p16); g32.zero_()
        self.p32, self.p16, self.g32 = p32, p16, g32
        self._bind_views()

    def sink(self, mode, flat_or_views, alpha: float = 1.0, remote=()) -> GradSink:
        views = flat_or_views if isinstance(flat_or_views, dict) else self._views(flat_or_views)
        return GradSink(mode, views, alpha, remote)

    def layer_bounds(self):
        """Flat ranges per transformer block (for bucketed gradient reduction)."""
        out = []
        for i in range(self.cfg.layers):
            es = [self.layout.entries[n] for n in self.layer_param_names(i).values()]
            lo = min(e.offset for e in es)
            hi = max(e.offset + -(-e.numel // 64) * 64 for e in es)
            out.append((lo, hi))
        return out

    def load_params(self, values: Dict[str, torch.Tensor]) -> None:
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

    # ---------------------------------------------------------------- helpers
    def _heads(self, t: torch.Tensor, which: Optional[int]) -> torch.Tensor:
        """[R, 3d] (which in 0..2) or [R, d] (which None) -> [B, H, T, hd] strided view."""
        c = self.cfg
        if which is None:
            return t.view(self.B, self.T, c.heads, c.head_dim).permute(0, 2, 1, 3)
        return t.view(self.B, self.T, 3, c.heads, c.head_dim)[:, :, which].permute(0, 2, 1, 3)

    # ---------------------------------------------------------------- forward
    def forward(self, ids: torch.Tensor, labels: torch.Tensor) -> None:
        c, W, W32 = self.cfg, self.w16, self.w32
        scale = 1.0 / math.sqrt(c.head_dim)
        ids = ids.reshape(-1).contiguous()
        post, ln = self.post_ln, self.norm_name()
        if post:
            ops.embedding_fwd(ids, W["model/wte"], W["model/wpe"], self.hf, self.T)
            ops.layernorm_fwd(self.hf, W32[ln + "/g"], W32[ln + "/b"], self.x[0], self.stat_f[0],
                              self.stat_f[1])
        else:
            ops.embedding_fwd(ids, W["model/wte"], W["model/wpe"], self.x[0], self.T)
        for i in range(c.layers):
            n = self.layer_param_names(i)
            m1, r1, m2, r2 = self.stats[i]
            if post:
                h_in = self.x[i]
            else:
                h_in = self.h1[i]
                ops.layernorm_fwd(self.x[i], W32[n["ln1_g"]], W32[n["ln1_b"]], h_in, m1, r1)
            ops.gemm(h_in, W[n["qkv_w"]], self.qkv[i], bias=W32[n["qkv_b"]])
            q, k, v = (self._heads(self.qkv[i], j) for j in range(3))
            if self.flash:
                # one kernel: row statistics, P (on chip only) and ctx = P V
                attn.fused_forward(q, k, v, self.ctx[i], self.att_m[i], self.att_il[i], scale,
                                   c.causal)
            else:
                if self.fused_attn:
                    attn.softmax_probs(q, k, self.P[i], self.att_stats, self.att_m[0],
                                       self.att_il[0], scale, c.causal)
                else:
                    ops.gemm_batched(q, k, self.S)
                    ops.attn_softmax_fwd(self.S, self.P[i], self.T, scale, c.causal)
                ops.gemm_batched(self.P[i], v, self._heads(self.ctx[i], None), b_mn=True)
            # x2 = x + proj(ctx): the residual stream (pre-LN) / LN1's input (post-LN)
            ops.gemm(self.ctx[i], W[n["o_w"]], self.x2[i], bias=W32[n["o_b"]], residual=self.x[i])
            if post:
                ops.layernorm_fwd(self.x2[i], W32[n["ln1_g"]], W32[n["ln1_b"]], self.h2[i], m1, r1)
            else:
                ops.layernorm_fwd(self.x2[i], W32[n["ln2_g"]], W32[n["ln2_b"]], self.h2[i], m2, r2)
            ops.gemm(self.h2[i], W[n["fc_w"]], self.f[i], bias=W32[n["fc_b"]], act=ops.ACT_GELU,
                     out2=self.z[i])
            if post:
                # h1 = h2 + mlp(h2) is LN2's input; its output is the next block's input
                ops.gemm(self.f[i], W[n["pr_w"]], self.h1[i], bias=W32[n["pr_b"]],
                         residual=self.h2[i])
                ops.layernorm_fwd(self.h1[i], W32[n["ln2_g"]], W32[n["ln2_b"]], self.x[i + 1], m2, r2)
            else:
                ops.gemm(self.f[i], W[n["pr_w"]], self.x[i + 1], bias=W32[n["pr_b"]],
                         residual=self.x2[i])
        if post:
            head_in = self.x[c.layers]
        else:
            head_in = self.hf
            ops.layernorm_fwd(self.x[c.layers], W32[ln + "/g"], W32[ln + "/b"], self.hf,
                              self.stat_f[0], self.stat_f[1])
        ops.gemm(head_in, W["model/wte"], self.logits)
        self.loss_buf.zero_()
        ops.softmax_xent_bf16_(self.logits, labels.reshape(-1).contiguous(), self.loss_buf,
                               c.vocab, 1.0 / self.R)

    # --------------------------------------------------------------- backward
    def _wgrad(self, dy, act, sink: GradSink, wname: str, bname: Optional[str]):
        ops.gemm(dy, act, sink.targets[wname], a_mn=True, b_mn=True, alpha=sink.alpha,
                 mode=sink.epi_for(wname))
        if bname is not None:
            ops.bias_grad(dy, sink.targets[bname], sink.alpha)

    def backward(self, ids: torch.Tensor, sink: GradSink) -> None:
        c, W, W32 = self.cfg, self.w16, self.w32
        ids = ids.reshape(-1).contiguous()
        if sink.mode == "store":
            # reduction-add targets (biases, LN params, embeddings) start from zero
            sink.zero_reduction_targets(extra=("model/wpe",))
        if self.post_ln:
            return self._backward_post_ln(ids, sink)
        da, db, dc = self.da
        # LM head (tied): dwte = dlogits^T hf ; dhf = dlogits . wte
        self._wgrad(self.dlogits, self.hf, sink, "model/wte", None)
        ops.gemm(self.dlogits, W["model/wte"], da, b_mn=True)
        ops.layernorm_bwd(da, self.x[c.layers], W32["model/ln_f/g"], self.stat_f[0], self.stat_f[1],
                          db, sink.targets["model/ln_f/g"], sink.targets["model/ln_f/b"], sink.alpha)
        dy = db                                  # gradient w.r.t. the block output x[i+1]
        spare = [da, dc]
        for i in range(c.layers - 1, -1, -1):
            n = self.layer_param_names(i)
            m1, r1, m2, r2 = self.stats[i]
            t0, t1 = spare
            # MLP
            self._wgrad(dy, self.f[i], sink, n["pr_w"], n["pr_b"])
            ops.gemm(dy, W[n["pr_w"]], self.dff, b_mn=True, mask=self.z[i], mask_mode=ops.MASK_GELU)
            self._wgrad(self.dff, self.h2[i], sink, n["fc_w"], n["fc_b"])
            ops.gemm(self.dff, W[n["fc_w"]], t0, b_mn=True)                      # dh2
            ops.layernorm_bwd(t0, self.x2[i], W32[n["ln2_g"]], m2, r2, t1, sink.targets[n["ln2_g"]],
                              sink.targets[n["ln2_b"]], sink.alpha, dx_add=dy)   # t1 = dx2
            # attention
            self._wgrad(t1, self.ctx[i], sink, n["o_w"], n["o_b"])
            ops.gemm(t1, W[n["o_w"]], t0, b_mn=True)                             # t0 = dctx
            self._attention_bwd(i, t0)                                           # dqkv
            self._wgrad(self.dqkv, self.h1[i], sink, n["qkv_w"], n["qkv_b"])
            ops.gemm(self.dqkv, W[n["qkv_w"]], t0, b_mn=True)                    # dh1
            ops.layernorm_bwd(t0, self.x[i], W32[n["ln1_g"]], m1, r1, dy, sink.targets[n["ln1_g"]],
                              sink.targets[n["ln1_b"]], sink.alpha, dx_add=t1)   # dy <- dx[i]
            if self.after_layer is not None:
                self.after_layer(i)
        ops.embedding_bwd(ids, dy, sink.targets["model/wte"], sink.targets["model/wpe"], self.T,
                          sink.alpha)

    def _backward_post_ln(self, ids: torch.Tensor, sink: GradSink) -> None:
        """BERT ordering.  Per block, with ``s1 = x + proj(ctx)``, ``h2 = LN1(s1)``,
        ``s2 = h2 + mlp(h2)``, ``x' = LN2(s2)``: the residual branches' gradients ride on the
        dgrad GEMMs' residual epilogue (``dh2 = dff . Wfc + ds2``, ``dx = dqkv . Wqkv + ds1``)."""
        c, W, W32 = self.cfg, self.w16, self.w32
        ln = self.norm_name()
        a, b, d_ = self.da
        L = c.layers
        self._wgrad(self.dlogits, self.x[L], sink, "model/wte", None)
        ops.gemm(self.dlogits, W["model/wte"], a, b_mn=True)          # a = dx[L]
        for i in range(L - 1, -1, -1):
            n = self.layer_param_names(i)
            m1, r1, m2, r2 = self.stats[i]
            ops.layernorm_bwd(a, self.h1[i], W32[n["ln2_g"]], m2, r2, b, sink.targets[n["ln2_g"]],
                              sink.targets[n["ln2_b"]], sink.alpha)              # b = ds2
            self._wgrad(b, self.f[i], sink, n["pr_w"], n["pr_b"])
            ops.gemm(b, W[n["pr_w"]], self.dff, b_mn=True, mask=self.z[i], mask_mode=ops.MASK_GELU)
            self._wgrad(self.dff, self.h2[i], sink, n["fc_w"], n["fc_b"])
            ops.gemm(self.dff, W[n["fc_w"]], d_, b_mn=True, residual=b)          # d_ = dh2
            ops.layernorm_bwd(d_, self.x2[i], W32[n["ln1_g"]], m1, r1, a, sink.targets[n["ln1_g"]],
                              sink.targets[n["ln1_b"]], sink.alpha)              # a = ds1
            self._wgrad(a, self.ctx[i], sink, n["o_w"], n["o_b"])
            ops.gemm(a, W[n["o_w"]], b, b_mn=True)                               # b = dctx
            self._attention_bwd(i, b)                                            # dqkv
            self._wgrad(self.dqkv, self.x[i], sink, n["qkv_w"], n["qkv_b"])
            ops.gemm(self.dqkv, W[n["qkv_w"]], d_, b_mn=True, residual=a)        # d_ = dx[i]
            a, d_ = d_, a
            if self.after_layer is not None:
                self.after_layer(i)
        ops.layernorm_bwd(a, self.hf, W32[ln + "/g"], self.stat_f[0], self.stat_f[1], b,
                          sink.targets[ln + "/g"], sink.targets[ln + "/b"], sink.alpha)
        ops.embedding_bwd(ids, b, sink.targets["model/wte"], sink.targets["model/wpe"], self.T,
                          sink.alpha)

    def _attention_bwd(self, i: int, dctx_flat: torch.Tensor) -> None:
        """``dctx`` ([R, d]) -> ``self.dqkv`` for block ``i``."""
        c = self.cfg
        scale = 1.0 / math.sqrt(c.head_dim)
        q, k, v = (self._heads(self.qkv[i], j) for j in range(3))
        dq, dk, dv = (self._heads(self.dqkv, j) for j in range(3))
        dctx = self._heads(dctx_flat, None)
        if self.flash:
            attn.rowdot(dctx_flat, self.ctx[i], self.att_D, self.B, self.T, c.heads, c.head_dim)
            attn.fused_backward(q, k, v, dk, dv, dctx, self.dS, self.att_m[i], self.att_il[i],
                                self.att_D, scale, c.causal)                     # dK, dV, dS
            ops.gemm_batched(self.dS, k, dq, b_mn=True)                          # dQ = dS K
            return
        ops.gemm_batched(self.P[i], dctx, dv, a_mn=True, b_mn=True)              # dV = P^T dctx
        if self.fused_attn:
            attn.rowdot(dctx_flat, self.ctx[i], self.att_D, self.B, self.T, c.heads, c.head_dim)
            attn.dscores(dctx, v, self.P[i], self.att_D, self.dS, scale, c.causal)
        else:
            ops.gemm_batched(dctx, v, self.S)                                    # dP = dctx V^T
            ops.attn_softmax_bwd(self.P[i], self.S, self.dS, self.T, scale)
        ops.gemm_batched(self.dS, k, dq, b_mn=True)                              # dQ = dS K
        ops.gemm_batched(self.dS, q, dk, a_mn=True, b_mn=True)                   # dK = dS^T Q

    def forward_backward_native(self, ids: torch.Tensor, labels: torch.Tensor,
                                sink: Optional[GradSink] = None) -> torch.Tensor:
        self.forward(ids, labels)
        self.backward(ids, sink or self.local_sink)
        return self.loss_buf

    def forward_backward(self, params, batch):
        self.load_params(params)
        ids, labels = batch
        loss = self.forward_backward_native(ids, labels).clone()
        return loss[0], self.export_grads()

    def local_apply(self, lr: float) -> None:
        ops.local_step(self.p32, self.g32, lr, p_bf16=self.p16)

    def train_step_local(self, ids, labels, lr: float) -> float:
        loss = self.forward_backward_native(ids, labels)
        self.local_apply(lr)
        return float(loss.item())

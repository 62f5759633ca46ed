"""Measured baseline: the same schemes written the way a careful
PyTorch user would -- ``torch`` ops (cuBLAS bf16 matmuls, cuDNN convolutions, flash SDPA,
autograd) for compute and NCCL collectives for every parameter-server interaction.
Nothing from ``dtg_b200`` is imported here.

Mapping of the ps traffic onto NCCL (the ps is sharded over all ranks, as in the native
arm's default layout):

* push of gradients to the ps shards     -> ``reduce_scatter_tensor`` (sum, fp32), one call
                                            per ~25 MB bucket, issued from an autograd hook the
                                            moment the bucket's last gradient exists, so it
                                            overlaps the rest of backward (what DDP does)
* ps-side SGD / Adagrad on fp32 master   -> one fused in-place op per bucket shard
* pull of parameters                      -> ONE ``all_gather_into_tensor`` of bf16 per bucket
                                            (fp32 for the windowed schemes, whose workers keep
                                            an fp32 local copy)

Gradients go from autograd straight into the flat fp32 communication buffer (one
converting copy per tensor, no intermediate bf16 flat buffer); with one rank there is no
communication and the hook applies the update to the fp32 master and refreshes the bf16
parameter in place.

For the asynchronous schemes this is the bulk-synchronous equivalent (every worker
contributes one gradient per round, like N Hogwild pushes between two pulls); NCCL has no
lock-free remote accumulate to express the real thing.
"""
from __future__ import annotations

import math
from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist
import torch.nn.functional as F

BUCKET_BYTES = 25 << 20          # DDP's default bucket size


class FlatState:
    """bf16 parameters (autograd leaves that are views of one flat buffer), fp32 master shards,
    fp32 gradient buffer, bucketed so that every bucket divides evenly over the ranks."""

    def __init__(self, shapes: Sequence[tuple], init: Callable[[int, tuple], torch.Tensor],
                 device, world: int, rank: int, adagrad: bool = False,
                 bucket_bytes: int = BUCKET_BYTES, keep_local32: bool = False):
        self.world, self.rank, self.dev = world, rank, device
        align = world * 8
        self.views, self.buckets = [], []          # (offset, numel, shape, bucket index)
        off, b_lo, cur = 0, 0, 0
        for i, s in enumerate(shapes):
            n = math.prod(s)
            self.views.append((off, n, tuple(s), len(self.buckets)))
            off += n
            cur += n * 4
            if cur >= bucket_bytes or i == len(shapes) - 1:
                off += (-(off - b_lo)) % align       # pad the bucket, not the tensors
                self.buckets.append((b_lo, off))
                b_lo, cur = off, 0
        self.total = off
        full = torch.zeros(self.total, device=device)
        for i, (o, n, s, _) in enumerate(self.views):
            full[o:o + n] = init(i, s).flatten().to(device)
        self.p16 = full.bfloat16()
        self.leaves = [self.p16[o:o + n].view(s).detach().requires_grad_(True)
                       for o, n, s, _ in self.views]
        self.g32 = torch.zeros(self.total, device=device) if world > 1 else None
        self.gviews = ([self.g32[o:o + n].view(s) for o, n, s, _ in self.views]
                       if world > 1 else None)
        self.master, self.acc, self.gshard = [], [], []
        for lo, hi in self.buckets:
            sh = (hi - lo) // world
            self.master.append(full[lo + rank * sh: lo + (rank + 1) * sh].clone())
            self.acc.append(torch.full((sh,), 0.1, device=device) if adagrad else None)
            self.gshard.append(torch.zeros(sh, device=device) if world > 1 else None)
        if world == 1:                    # per-tensor views of the (single-shard) master
            self.mviews = [self.master[b][o - self.buckets[b][0]: o - self.buckets[b][0] + n].view(s)
                           for o, n, s, b in self.views]
        self.local32 = full.clone() if keep_local32 else None
        self.grad_scale = 1.0             # per-worker learning-rate ratio (ssgd_lr), folded in here
        self._remaining = None
        self._works = {}

    # ------------------------------------------------------------------ hooks
    def arm(self, on_bucket_ready: Callable[[int], None]):
        per_bucket = [0] * len(self.buckets)
        for _, _, _, b in self.views:
            per_bucket[b] += 1
        self._per_bucket = per_bucket

        def make(i):
            b = self.views[i][3]

            def hook(p):
                # bf16 -> fp32, straight into the comm buffer
                if self.grad_scale == 1.0:
                    self.gviews[i].copy_(p.grad)
                else:
                    torch.mul(p.grad, self.grad_scale, out=self.gviews[i])
                p.grad = None
                self._remaining[b] -= 1
                if self._remaining[b] == 0:
                    on_bucket_ready(b)
            return hook
        for i, p in enumerate(self.leaves):
            p.register_post_accumulate_grad_hook(make(i))

    def begin_backward(self):
        self._remaining = list(self._per_bucket)

    # ------------------------------------------------------------------ comm
    def reduce_scatter_async(self, b: int, src: Optional[torch.Tensor] = None):
        lo, hi = self.buckets[b]
        buf = self.g32 if src is None else src
        self._works[b] = dist.reduce_scatter_tensor(self.gshard[b], buf[lo:hi],
                                                    op=dist.ReduceOp.SUM, async_op=True)

    def apply_and_gather(self, lr: float, scale: float, adagrad: bool, gather_fp32: bool = False):
        """wait(bucket) -> fused update of the master shard -> all-gather, bucket by bucket in
        the order the reductions were issued; returns when every parameter is refreshed."""
        ags = []
        for b in sorted(self._works, key=lambda k: -k):
            self._works[b].wait()
            g, m = self.gshard[b], self.master[b]
            if adagrad:
                a = self.acc[b]
                a.addcmul_(g, g, value=scale * scale)
                m.addcdiv_(g, a.sqrt(), value=-lr * scale)
            else:
                m.add_(g, alpha=-lr * scale)
            lo, hi = self.buckets[b]
            if gather_fp32:
                ags.append(dist.all_gather_into_tensor(self.local32[lo:hi], m, async_op=True))
            else:
                ags.append(dist.all_gather_into_tensor(self.p16[lo:hi], m.bfloat16(), async_op=True))
        for w in ags:
            w.wait()
        self._works.clear()
        if gather_fp32:
            self.p16.copy_(self.local32)


class TorchBaseline:
    """MLP / ResNet-50 (``forward`` closure) under every scheme."""

    def __init__(self, sizes: List[int], batch: int, device, scheme: str = "hogwild",
                 lr: float = 1e-3, update_window: int = 8, world: int = 1, rank: int = 0,
                 shapes=None, forward=None, ones=(), loss_fn=None, init_std=None,
                 worker_lrs: Optional[Sequence[float]] = None):
        self.sizes, self.B, self.dev = sizes, batch, device
        self.scheme, self.lr, self.T = scheme, lr, update_window
        self.world, self.rank = world, rank
        self._forward, self._loss_fn = forward, loss_fn
        g = torch.Generator().manual_seed(0)
        if shapes is None:
            shapes = []
            for i in range(len(sizes) - 1):
                shapes += [(sizes[i + 1], sizes[i]), (sizes[i + 1],)]
        self.shapes = shapes
        ones = set(ones)

        def init(i, s):
            if init_std is not None:
                return init_std(i, s, g)
            if len(s) >= 2:
                fan = math.prod(s[1:]) + s[0]
                return (torch.rand(s, generator=g) * 2 - 1) * math.sqrt(6.0 / fan)
            return torch.ones(s) if i in ones else torch.zeros(s)
        self.windowed = scheme in ("downpour", "downpour_easy", "adag", "sdag")
        self.ps_adagrad = scheme in ("downpour", "downpour_easy")
        self.st = FlatState(shapes, init, device, world, rank, adagrad=self.ps_adagrad,
                            keep_local32=self.windowed)
        st = self.st
        if self.windowed:
            self.window = torch.zeros(st.total, device=device)
            self.gflat = torch.zeros(st.total, device=device, dtype=torch.bfloat16)
            self.gfviews = [self.gflat[o:o + n].view(s) for o, n, s, _ in st.views]
            self.local_acc = torch.full((st.total,), 0.1, device=device) if scheme == "downpour" else None
        elif world > 1:
            st.arm(lambda b: st.reduce_scatter_async(b))
            if scheme == "ssgd_lr" and worker_lrs is not None:
                # Synchronous-SGD-different-learning-rates/ssgd.py:59-64: w -= sum_k lr_k g_k
                st.grad_scale = worker_lrs[rank] / lr
        else:
            # one rank: the "ps" is local; apply the update the moment a gradient exists
            def make(i):
                def hook(p):
                    st.mviews[i].add_(p.grad, alpha=-self.lr)
                    p.data.copy_(st.mviews[i])
                    p.grad = None
                return hook
            for i, p in enumerate(st.leaves):
                p.register_post_accumulate_grad_hook(make(i))

    # ------------------------------------------------------------------
    def _loss(self, ps, x, y):
        if self._forward is not None:
            h = self._forward(ps, x)
        else:
            h = x.bfloat16()
            L = len(ps) // 2
            for i in range(L):
                h = F.linear(h, ps[2 * i], ps[2 * i + 1])
                if i + 1 < L:
                    h = F.relu(h)
        if self._loss_fn is not None:
            return self._loss_fn(h, y)
        return F.cross_entropy(h.float(), y)

    def step(self, x, y):
        s, st = self.scheme, self.st
        if not self.windowed:
            if self.world > 1:
                st.begin_backward()
            loss = self._loss(st.leaves, x, y)
            loss.backward()
            if self.world > 1:
                # hogwild: sum of pushes; ssgd_lr: sum of lr-scaled gradients; others: mean
                scale = 1.0 if (s == "hogwild" or st.grad_scale != 1.0 or
                                (s == "ssgd_lr" and self.world > 1)) else 1.0 / self.world
                st.apply_and_gather(self.lr, scale, False)
            return loss.detach()
        # windowed schemes: T local steps on an fp32 local copy, then push the window.
        # x / y may hold T micro-batches stacked along dim 0 (fresh data for every local step).
        T = self.T
        stacked = x.shape[0] == T * self.B and T > 1
        xs = x.chunk(T) if stacked else [x] * T
        ys = y.chunk(T) if stacked else [y] * T
        self.window.zero_()
        loss = None
        for xb, yb in zip(xs, ys):
            loss = self._loss(st.leaves, xb, yb)
            grads = torch.autograd.grad(loss, st.leaves)
            torch._foreach_copy_(self.gfviews, list(grads))        # one multi-tensor launch
            gf = self.gflat
            self.window.add_(gf)
            if s == "downpour":
                self.local_acc.addcmul_(gf, gf)
                st.local32.addcdiv_(gf, self.local_acc.sqrt(), value=-self.lr)
            else:
                st.local32.add_(gf, alpha=-self.lr)
            st.p16.copy_(st.local32)
        scale = {"adag": 1.0 / T, "sdag": 1.0 / (T * self.world)}.get(s, 1.0)
        if self.world > 1:
            for b in reversed(range(len(st.buckets))):
                st.reduce_scatter_async(b, self.window)
            st.apply_and_gather(self.lr, scale, self.ps_adagrad, gather_fp32=True)
        else:
            for b, (lo, hi) in enumerate(st.buckets):
                m, w = st.master[b], self.window[lo:hi]
                if self.ps_adagrad:
                    st.acc[b].addcmul_(w, w, value=scale * scale)
                    m.addcdiv_(w, st.acc[b].sqrt(), value=-self.lr * scale)
                else:
                    m.add_(w, alpha=-self.lr * scale)
                st.local32[lo:hi].copy_(m)
            st.p16.copy_(st.local32)
        return loss.detach()

    def flat_params(self) -> torch.Tensor:
        """Current bf16 parameters without bucket padding (loss-parity checks)."""
        return torch.cat([self.st.p16[o:o + n] for o, n, _, _ in self.st.views])


class TorchTransformerBaseline(TorchBaseline):
    """GPT-2 medium / BERT-large dims the PyTorch way: bf16 matmuls (cuBLAS),
    ``F.scaled_dot_product_attention`` (flash), autograd, fp32 master weights, bucketed NCCL
    reduce-scatter overlapped with backward + sharded SGD + bf16 all-gather."""

    def __init__(self, cfg, batch: int, device, scheme: str = "ssgd", lr: float = 1e-3,
                 world: int = 1, rank: int = 0, worker_lrs=None):
        c, d = cfg, cfg.dim
        self.cfg = cfg
        post = bool(getattr(cfg, "post_ln", False))
        spec = [("wte", (c.vocab, d), 0.02), ("wpe", (c.seq, d), 0.01)]
        norm = [("lnfg", (d,), 1), ("lnfb", (d,), 0)]       # final LN (pre-LN) / embedding LN (post-LN)
        if post:
            spec += norm
        for i in range(c.layers):
            spec += [("ln1g%d" % i, (d,), 1), ("ln1b%d" % i, (d,), 0),
                     ("qkvw%d" % i, (3 * d, d), 0.02), ("qkvb%d" % i, (3 * d,), 0),
                     ("ow%d" % i, (d, d), 0.02), ("ob%d" % i, (d,), 0),
                     ("ln2g%d" % i, (d,), 1), ("ln2b%d" % i, (d,), 0),
                     ("fcw%d" % i, (c.ff, d), 0.02), ("fcb%d" % i, (c.ff,), 0),
                     ("prw%d" % i, (d, c.ff), 0.02), ("prb%d" % i, (d,), 0)]
        if not post:
            spec += norm
        self.index = {name: i for i, (name, _, _) in enumerate(spec)}

        def init_std(i, s, g):
            v = spec[i][2]
            if len(s) == 2:
                return torch.randn(s, generator=g) * v
            return torch.ones(s) if v == 1 else torch.zeros(s)

        def forward(ps, ids):
            B, T = ids.shape
            P = lambda n: ps[self.index[n]]   # noqa: E731
            ln = lambda t, g, b: F.layer_norm(t, (c.dim,), P(g), P(b), 1e-5)   # noqa: E731
            x = P("wte")[ids] + P("wpe")[:T][None]
            if post:
                x = ln(x, "lnfg", "lnfb")
            for i in range(c.layers):
                h = x if post else ln(x, "ln1g%d" % i, "ln1b%d" % i)
                qkv = F.linear(h, P("qkvw%d" % i), P("qkvb%d" % i))
                q, k, v = qkv.view(B, T, 3, c.heads, c.head_dim).permute(2, 0, 3, 1, 4)
                ctx = F.scaled_dot_product_attention(q, k, v, is_causal=c.causal)
                ctx = ctx.transpose(1, 2).reshape(B, T, c.dim)
                x = x + F.linear(ctx, P("ow%d" % i), P("ob%d" % i))
                if post:
                    x = ln(x, "ln1g%d" % i, "ln1b%d" % i)
                h = x if post else ln(x, "ln2g%d" % i, "ln2b%d" % i)
                f = F.gelu(F.linear(h, P("fcw%d" % i), P("fcb%d" % i)), approximate="tanh")
                x = x + F.linear(f, P("prw%d" % i), P("prb%d" % i))
                if post:
                    x = ln(x, "ln2g%d" % i, "ln2b%d" % i)
            if not post:
                x = ln(x, "lnfg", "lnfb")
            return F.linear(x, P("wte"))

        def loss_fn(logits, labels):
            # bf16 logits straight into the fused loss (no fp32 copy of the [B*T, vocab] tensor)
            return F.cross_entropy(logits.view(-1, c.vocab), labels.reshape(-1))
        super().__init__([], batch, device, scheme=scheme, lr=lr, world=world, rank=rank,
                         shapes=[s for _, s, _ in spec], forward=forward, loss_fn=loss_fn,
                         init_std=init_std, worker_lrs=worker_lrs)


def torch_resnet50_baseline(batch: int, device, **kw) -> TorchBaseline:
    """ResNet-50 (v1.5) the PyTorch way: channels_last bf16 ``F.conv2d`` (cuDNN),
    ``F.batch_norm`` with batch statistics, autograd; ps traffic over NCCL as above."""
    blocks, widths = (3, 4, 6, 3), (64, 128, 256, 512)
    plan = []
    cin = 64
    for b, (n, w) in enumerate(zip(blocks, widths)):
        for u in range(n):
            stride = 2 if (u == 0 and b > 0) else 1
            unit = [(cin, w, 1, 1, 0), (w, w, 3, stride, 1), (w, 4 * w, 1, 1, 0)]
            sc = (cin, 4 * w, 1, stride, 0) if (cin != 4 * w or stride != 1) else None
            plan.append((unit, sc))
            cin = 4 * w
    shapes, ones = [], []

    def add_conv(ci, co, k):
        shapes.append((co, ci, k, k)); ones.append(len(shapes)); shapes.append((co,)); shapes.append((co,))
    add_conv(3, 64, 7)
    for unit, sc in plan:
        for ci, co, k, s_, p_ in unit:
            add_conv(ci, co, k)
        if sc:
            add_conv(sc[0], sc[1], 1)
    shapes += [(1000, 2048), (1000,)]

    def conv_bn(ps, i, x, stride, pad, relu, res=None):
        y = F.conv2d(x, ps[i], None, stride, pad)
        y = F.batch_norm(y, None, None, ps[i + 1], ps[i + 2], True, 0.0, 1e-5)
        if res is not None:
            y = y + res
        return F.relu(y) if relu else y

    def forward(ps, x):
        h = x.bfloat16().permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
        i = 0
        h = conv_bn(ps, i, h, 2, 3, True); i += 3
        h = F.max_pool2d(h, 3, 2, 1)
        for unit, sc in plan:
            j = i + 9
            idn = h
            if sc:
                idn = conv_bn(ps, j, h, sc[3], 0, False)
            a = conv_bn(ps, i, h, 1, 0, True)
            b = conv_bn(ps, i + 3, a, unit[1][3], 1, True)
            h = conv_bn(ps, i + 6, b, 1, 0, True, res=idn)
            i = j + (3 if sc else 0)
        h = h.mean((2, 3))
        return F.linear(h, ps[i], ps[i + 1])

    return TorchBaseline([], batch, device, shapes=shapes, forward=forward, ones=ones, **kw)

"""Measured baseline: the same schemes written the way a PyTorch
user would -- ``torch`` ops (cuBLAS bf16 matmuls, autograd) for compute and NCCL
collectives for every parameter-server interaction.  Nothing from ``dtg_b200`` is
imported here.

Mapping of the ps traffic onto NCCL (the ps is sharded over all ranks, as in the
native arm's default layout):

* push of gradients to the ps shards     -> ``reduce_scatter_tensor`` (sum)
* ps-side SGD / Adagrad on fp32 master   -> elementwise torch ops on the shard
* pull of parameters                      -> ``all_gather_into_tensor``

For the asynchronous schemes this is the bulk-synchronous equivalent (every worker
contributes one gradient per round, like N Hogwild pushes between two pulls); NCCL
has no lock-free remote accumulate to express the real thing.
"""
from __future__ import annotations

import math
from typing import List

import torch
import torch.distributed as dist
import torch.nn.functional as F


class TorchBaseline:
    def __init__(self, sizes: List[int], batch: int, device, scheme: str = "hogwild",
                 lr: float = 1e-3, update_window: int = 8, world: int = 1, rank: int = 0,
                 shapes=None, forward=None, ones=()):
        self.sizes, self.B, self.dev = sizes, batch, device
        self.scheme, self.lr, self.T = scheme, lr, update_window
        self.world, self.rank = world, rank
        self._forward = forward
        g = torch.Generator().manual_seed(0)
        if shapes is None:
            shapes = []
            for i in range(len(sizes) - 1):
                shapes += [(sizes[i + 1], sizes[i]), (sizes[i + 1],)]
        self.shapes = shapes
        numels = [math.prod(s) for s in shapes]
        total = sum(numels)
        pad = (-total) % (world * 4)
        self.total = total + pad
        self.master = torch.zeros(self.total, device=device)            # fp32, full copy
        off = 0
        for s, n in zip(shapes, numels):
            if len(s) >= 2:
                fan = math.prod(s[1:]) + s[0]
                lim = math.sqrt(6.0 / fan)
                self.master[off:off + n] = ((torch.rand(s, generator=g) * 2 - 1) * lim).flatten().to(device)
            off += n
        for i in ones:
            o = sum(numels[:i])
            self.master[o:o + numels[i]] = 1.0
        self.p16 = self.master.bfloat16()
        self.p16.requires_grad_(False)
        self.views = []
        off = 0
        for s, n in zip(shapes, numels):
            self.views.append((off, n, s))
            off += n
        self.shard = self.total // world
        self.g32 = torch.zeros(self.total, device=device)
        self.gshard = torch.zeros(self.shard, device=device)
        self.acc = torch.full((self.shard,), 0.1, device=device)
        self.window = torch.zeros(self.total, device=device)
        self.local32 = self.master.clone()
        self.local_acc = torch.full((self.total,), 0.1, device=device)
        self._gflat16 = torch.zeros(self.total, device=device, dtype=torch.bfloat16)

    def _params(self, flat16):
        return [flat16[o:o + n].view(s) for o, n, s in self.views]

    def _fwd_bwd(self, flat16, x, y):
        # parameters are views of the flat bf16 buffer, each its own autograd leaf; their
        # gradients are written back into one flat buffer (what DDP-style bucketing does)
        ps = [p.detach().requires_grad_(True) for p in self._params(flat16)]
        if self._forward is not None:
            h = self._forward(ps, x)
        else:
            h = x.bfloat16()
            L = len(ps) // 2
            for i in range(L):
                h = F.linear(h, ps[2 * i], ps[2 * i + 1])
                if i + 1 < L:
                    h = F.relu(h)
        loss = F.cross_entropy(h.float(), y)
        grads = torch.autograd.grad(loss, ps)
        g = self._gflat16
        for (o, n, _), gi in zip(self.views, grads):
            g[o:o + n].copy_(gi.reshape(-1))
        return loss.detach(), g

    def _ps_apply(self, gflat32, adagrad: bool, scale: float = 1.0):
        """reduce_scatter -> sharded update of the fp32 master -> all_gather (bf16 + fp32)."""
        sl = slice(self.rank * self.shard, (self.rank + 1) * self.shard)
        if self.world > 1:
            dist.reduce_scatter_tensor(self.gshard, gflat32, op=dist.ReduceOp.SUM)
            gs = self.gshard
        else:
            gs = gflat32
        if scale != 1.0:
            gs = gs * scale
        m = self.master[sl]
        if adagrad:
            self.acc.addcmul_(gs, gs)
            m.addcdiv_(gs, self.acc.sqrt(), value=-self.lr)
        else:
            m.add_(gs, alpha=-self.lr)
        if self.world > 1:
            dist.all_gather_into_tensor(self.p16, m.bfloat16())
            dist.all_gather_into_tensor(self.local32, m.contiguous())
        else:
            self.p16.copy_(m)
            self.local32.copy_(m)

    def step(self, x, y):
        s = self.scheme
        if s in ("hogwild", "ssgd", "ssgd_lr", "tower_avg", "multigpu_async"):
            loss, g = self._fwd_bwd(self.p16, x, y)
            self.g32.copy_(g)
            scale = 1.0 if s == "hogwild" else 1.0 / self.world      # sum of pushes vs mean
            self._ps_apply(self.g32, False, scale)
            return loss
        # windowed schemes: T local steps on a local fp32 copy, then push the window
        self.window.zero_()
        local16 = self.p16.clone()
        loss = None
        for _ in range(self.T):
            loss, g = self._fwd_bwd(local16, x, y)
            g32 = g.float()
            self.window.add_(g32)
            if s == "downpour":
                self.local_acc.addcmul_(g32, g32)
                self.local32.addcdiv_(g32, self.local_acc.sqrt(), value=-self.lr)
            else:
                self.local32.add_(g32, alpha=-self.lr)
            local16.copy_(self.local32)
        if s == "adag":
            self._ps_apply(self.window, False, 1.0 / self.T)
        elif s == "sdag":
            self._ps_apply(self.window, False, 1.0 / (self.T * self.world))
        else:
            self._ps_apply(self.window, True)
        return loss


class TorchTransformerBaseline:
    """GPT-2 medium / BERT-large the PyTorch way: bf16 autocast-style matmuls (cuBLAS),
    ``F.scaled_dot_product_attention`` (flash), autograd, fp32 master weights, NCCL
    reduce-scatter + sharded SGD + all-gather for the (sharded) parameter server."""

    def __init__(self, cfg, batch: int, device, scheme: str = "ssgd", lr: float = 1e-3,
                 world: int = 1, rank: int = 0):
        self.cfg, self.B, self.dev = cfg, batch, device
        self.scheme, self.lr, self.world, self.rank = scheme, lr, world, rank
        c, d = cfg, cfg.dim
        g = torch.Generator().manual_seed(0)
        shapes = [("wte", (c.vocab, d), 0.02), ("wpe", (c.seq, d), 0.01)]
        for i in range(c.layers):
            shapes += [("ln1g%d" % i, (d,), 1), ("ln1b%d" % i, (d,), 0),
                       ("qkvw%d" % i, (3 * d, d), 0.02), ("qkvb%d" % i, (3 * d,), 0),
                       ("ow%d" % i, (d, d), 0.02), ("ob%d" % i, (d,), 0),
                       ("ln2g%d" % i, (d,), 1), ("ln2b%d" % i, (d,), 0),
                       ("fcw%d" % i, (c.ff, d), 0.02), ("fcb%d" % i, (c.ff,), 0),
                       ("prw%d" % i, (d, c.ff), 0.02), ("prb%d" % i, (d,), 0)]
        shapes += [("lnfg", (d,), 1), ("lnfb", (d,), 0)]
        self.views, off = {}, 0
        for name, shp, _ in shapes:
            n = math.prod(shp)
            self.views[name] = (off, n, shp)
            off += n
        self.total = off + ((-off) % (world * 4))
        self.master = torch.zeros(self.total, device=device)
        for name, shp, init in shapes:
            o, n, _ = self.views[name]
            if len(shp) == 2:
                self.master[o:o + n] = (torch.randn(shp, generator=g) * init).flatten().to(device)
            elif init == 1:
                self.master[o:o + n] = 1.0
        self.p16 = self.master.bfloat16()
        self.shard = self.total // world
        self.g32 = torch.zeros(self.total, device=device)
        self.gshard = torch.zeros(self.shard, device=device)

    def _p(self, flat, name):
        o, n, shp = self.views[name]
        return flat[o:o + n].view(shp)

    def _loss(self, leaves, ids, labels):
        c = self.cfg
        B, T = ids.shape
        P = lambda n: leaves[n]   # noqa: E731
        x = P("wte")[ids] + P("wpe")[:T][None]
        for i in range(c.layers):
            h = F.layer_norm(x, (c.dim,), P("ln1g%d" % i), P("ln1b%d" % i), 1e-5)
            qkv = F.linear(h, P("qkvw%d" % i), P("qkvb%d" % i))
            q, k, v = qkv.view(B, T, 3, c.heads, c.head_dim).permute(2, 0, 3, 1, 4)
            ctx = F.scaled_dot_product_attention(q, k, v, is_causal=c.causal)
            ctx = ctx.transpose(1, 2).reshape(B, T, c.dim)
            x = x + F.linear(ctx, P("ow%d" % i), P("ob%d" % i))
            h = F.layer_norm(x, (c.dim,), P("ln2g%d" % i), P("ln2b%d" % i), 1e-5)
            f = F.gelu(F.linear(h, P("fcw%d" % i), P("fcb%d" % i)), approximate="tanh")
            x = x + F.linear(f, P("prw%d" % i), P("prb%d" % i))
        x = F.layer_norm(x, (c.dim,), P("lnfg"), P("lnfb"), 1e-5)
        logits = F.linear(x, P("wte"))
        return F.cross_entropy(logits.float().view(-1, c.vocab), labels.reshape(-1))

    def step(self, ids, labels):
        names = list(self.views)
        leaves = {n: self._p(self.p16, n).detach().requires_grad_(True) for n in names}
        loss = self._loss(leaves, ids, labels)
        grads = torch.autograd.grad(loss, [leaves[n] for n in names])
        for n, gi in zip(names, grads):
            o, cnt, _ = self.views[n]
            self.g32[o:o + cnt].copy_(gi.reshape(-1))
        sl = slice(self.rank * self.shard, (self.rank + 1) * self.shard)
        if self.world > 1:
            dist.reduce_scatter_tensor(self.gshard, self.g32, op=dist.ReduceOp.SUM)
            gs = self.gshard / self.world
        else:
            gs = self.g32
        m = self.master[sl]
        m.add_(gs, alpha=-self.lr)
        if self.world > 1:
            dist.all_gather_into_tensor(self.p16, m.bfloat16())
        else:
            self.p16.copy_(m)
        return loss.detach()


def torch_resnet50_baseline(batch: int, device, **kw) -> TorchBaseline:
    """ResNet-50 (v1.5) the PyTorch way: channels_last bf16 ``F.conv2d`` (cuDNN),
    ``F.batch_norm`` with batch statistics, autograd; ps traffic over NCCL as above."""
    blocks, widths = (3, 4, 6, 3), (64, 128, 256, 512)
    convs = [("stem", 3, 64, 7, 2, 3)]
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

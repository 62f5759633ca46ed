"""ResNet (v1.5 bottleneck) for BASELINE.json config 3: DOWNPOUR / ADAG with Adagrad on
ResNet-50, synthetic 224x224 images, batch 256 per worker.  Not a reference model (the
reference's only model is the two-vector toy): it is the workload the north star names.

Parameter names follow TF-slim (``resnet_v1_50/block2/unit_1/bottleneck_v1/conv2/weights``,
``.../BatchNorm/gamma``).  Convolution weights are stored ``[Cout, kh, kw, Cin]`` flattened to
``[Cout, kh*kw*Cin]`` -- the K-major B operand of the im2col GEMM; the stem's 3 input
channels are padded to 8 (zero weights) so every row pitch is a multiple of 16 bytes.
BatchNorm uses batch statistics (training mode); moving averages are not tracked.

Plain-PyTorch reference implementation (NCHW ``F.conv2d``); ``resnet_native.py`` is the
sm_100a path.
"""
from __future__ import annotations

import dataclasses
import math
from typing import List, Tuple

import torch
import torch.nn.functional as F

from .base import Model, ParamSpec

__all__ = ["ResNetConfig", "ResNet", "resnet50", "tiny_resnet", "ConvSpec"]

CIN_PAD = 8


@dataclasses.dataclass
class ResNetConfig:
    blocks: Tuple[int, ...] = (3, 4, 6, 3)
    widths: Tuple[int, ...] = (64, 128, 256, 512)
    stem: int = 64
    classes: int = 1000
    image: int = 224
    name: str = "resnet_v1_50"


@dataclasses.dataclass
class ConvSpec:
    name: str          # scope, e.g. resnet_v1_50/conv1
    cin: int
    cout: int
    k: int
    stride: int
    pad: int
    relu: bool


def _he(shape, gen):
    fan_in = shape[1]
    return torch.randn(shape, generator=gen) * math.sqrt(2.0 / fan_in)


class ResNet(Model):
    def __init__(self, cfg: ResNetConfig, compute_dtype=torch.float32):
        self.cfg, self.name, self.compute_dtype = cfg, cfg.name, compute_dtype

    # ---- structure ---------------------------------------------------------
    def stem_conv(self) -> ConvSpec:
        return ConvSpec(self.cfg.name + "/conv1", CIN_PAD, self.cfg.stem, 7, 2, 3, True)

    def units(self):
        """Yields (scope, [conv1, conv2, conv3], shortcut_or_None)."""
        c = self.cfg
        cin = c.stem
        for b, (n, w) in enumerate(zip(c.blocks, c.widths), 1):
            for u in range(1, n + 1):
                stride = 2 if (u == 1 and b > 1) else 1
                scope = "%s/block%d/unit_%d/bottleneck_v1" % (c.name, b, u)
                convs = [ConvSpec(scope + "/conv1", cin, w, 1, 1, 0, True),
                         ConvSpec(scope + "/conv2", w, w, 3, stride, 1, True),
                         ConvSpec(scope + "/conv3", w, 4 * w, 1, 1, 0, False)]
                sc = ConvSpec(scope + "/shortcut", cin, 4 * w, 1, stride, 0, False) \
                    if (cin != 4 * w or stride != 1) else None
                yield scope, convs, sc
                cin = 4 * w

    def all_convs(self) -> List[ConvSpec]:
        out = [self.stem_conv()]
        for _, convs, sc in self.units():
            out += convs + ([sc] if sc else [])
        return out

    def param_specs(self) -> List[ParamSpec]:
        specs = []
        for cv in self.all_convs():
            fan_in = cv.k * cv.k * cv.cin
            specs.append(ParamSpec(cv.name + "/weights", (cv.cout, fan_in), _he,
                                   ("normal", math.sqrt(2.0 / fan_in))))
            specs.append(ParamSpec(cv.name + "/BatchNorm/gamma", (cv.cout,),
                                   lambda s, g: torch.ones(s), ("constant", 1.0)))
            specs.append(ParamSpec(cv.name + "/BatchNorm/beta", (cv.cout,),
                                   lambda s, g: torch.zeros(s), ("constant", 0.0)))
        feat = 4 * self.cfg.widths[-1]
        specs.append(ParamSpec(self.cfg.name + "/logits/weights", (self.cfg.classes, feat),
                               lambda s, g: torch.randn(s, generator=g) * 0.01, ("normal", 0.01)))
        specs.append(ParamSpec(self.cfg.name + "/logits/biases", (self.cfg.classes,),
                               lambda s, g: torch.zeros(s), ("constant", 0.0)))
        return specs

    # ---- reference forward (NCHW) -------------------------------------------
    def _conv_bn(self, P, cv: ConvSpec, x, residual=None):
        w = P[cv.name + "/weights"].view(cv.cout, cv.k, cv.k, cv.cin).permute(0, 3, 1, 2)
        y = F.conv2d(x, w, None, cv.stride, cv.pad)
        y = F.batch_norm(y, None, None, P[cv.name + "/BatchNorm/gamma"],
                         P[cv.name + "/BatchNorm/beta"], True, 0.0, 1e-5)
        if residual is not None:
            y = y + residual
        return F.relu(y) if (cv.relu or residual is not None) else y

    def logits(self, params, x):
        """``x``: [N, H, W, 8] channels-last (3 real channels + zero padding)."""
        dt = self.compute_dtype
        P = {k: v.to(dt) for k, v in params.items()}
        h = x.to(dt).permute(0, 3, 1, 2)
        h = self._conv_bn(P, self.stem_conv(), h)
        h = F.max_pool2d(h, 3, 2, 1)
        for _, convs, sc in self.units():
            idn = h if sc is None else self._conv_bn(P, sc, h)
            a = self._conv_bn(P, convs[0], h)
            b = self._conv_bn(P, convs[1], a)
            h = self._conv_bn(P, convs[2], b, residual=idn)
        h = h.mean((2, 3))
        return F.linear(h, P[self.cfg.name + "/logits/weights"], P[self.cfg.name + "/logits/biases"])

    def loss(self, params, batch):
        x, y = batch
        return F.cross_entropy(self.logits(params, x).float(), y)

    def synthetic_batch(self, batch_size, device="cpu", generator=None):
        s = self.cfg.image
        x = torch.zeros(batch_size, s, s, CIN_PAD)
        x[..., :3] = torch.randn(batch_size, s, s, 3, generator=generator)
        y = torch.randint(0, self.cfg.classes, (batch_size,), generator=generator)
        return x.to(device), y.to(device)

    def flops_per_sample(self) -> float:
        s = self.cfg.image
        total, h = 0.0, s
        sizes = {}
        h = (s + 6 - 7) // 2 + 1
        total += 2.0 * h * h * 49 * 3 * self.cfg.stem
        h = (h + 2 - 3) // 2 + 1
        for _, convs, sc in self.units():
            hin = h
            for cv in convs:
                ho = (hin + 2 * cv.pad - cv.k) // cv.stride + 1
                total += 2.0 * ho * ho * cv.k * cv.k * cv.cin * cv.cout
                hin = ho
            if sc:
                total += 2.0 * hin * hin * sc.cin * sc.cout
            h = hin
        return 3.0 * total


def resnet50(**kw) -> ResNet:
    """ResNet-50: ~25.6 M parameters (BASELINE.json config 3)."""
    return ResNet(ResNetConfig(), **kw)


def tiny_resnet(**kw) -> ResNet:
    return ResNet(ResNetConfig((1, 2), (16, 32), 16, 10, 32, "resnet_tiny"), **kw)

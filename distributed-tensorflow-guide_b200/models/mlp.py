"""Fully connected classifiers from BASELINE.json: the MNIST-shape MLP
784-100-10 (config 1, CPU/gloo plumbing) and the wide MLP 784-4096x4-10
(config 2, Hogwild on 8xB200).  Parameter names follow ``tf.layers.dense``:
``dense/kernel``, ``dense/bias``, ``dense_1/kernel`` ...  Kernels are stored
``[out, in]`` (K-major for the tcgen05 GEMMs); hidden activations are ReLU; the
loss is mean softmax cross-entropy.

This file is the plain-PyTorch reference implementation (used on CPU and as the
torch/cuBLAS baseline); ``models/mlp_native.py`` is the sm_100a path.
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence

import torch
import torch.nn.functional as F

from .base import Model, ParamSpec

__all__ = ["MLP", "mnist_mlp", "wide_mlp"]


def _glorot(shape, gen):
    fan_out, fan_in = shape
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return (torch.rand(shape, generator=gen) * 2 - 1) * lim


class MLP(Model):
    name = "mlp"

    def __init__(self, sizes: Sequence[int], compute_dtype: torch.dtype = torch.float32):
        self.sizes = list(sizes)
        self.compute_dtype = compute_dtype

    @property
    def num_layers(self) -> int:
        return len(self.sizes) - 1

    def layer_names(self, i: int):
        p = "dense" if i == 0 else "dense_%d" % i
        return p + "/kernel", p + "/bias"

    def param_specs(self) -> List[ParamSpec]:
        out = []
        for i in range(self.num_layers):
            k, b = self.layer_names(i)
            lim = math.sqrt(6.0 / (self.sizes[i] + self.sizes[i + 1]))
            out.append(ParamSpec(k, (self.sizes[i + 1], self.sizes[i]), _glorot, ("uniform", lim)))
            out.append(ParamSpec(b, (self.sizes[i + 1],), lambda s, g: torch.zeros(s),
                                 ("constant", 0.0)))
        return out

    def logits(self, params: Dict[str, torch.Tensor], x: torch.Tensor) -> torch.Tensor:
        dt = self.compute_dtype
        h = x.to(dt)
        for i in range(self.num_layers):
            k, b = self.layer_names(i)
            h = F.linear(h, params[k].to(dt), params[b].to(dt))
            if i + 1 < self.num_layers:
                h = F.relu(h)
        return h

    def loss(self, params, batch):
        x, y = batch
        return F.cross_entropy(self.logits(params, x).float(), y)

    def synthetic_batch(self, batch_size, device="cpu", generator=None):
        x = torch.randn(batch_size, self.sizes[0], generator=generator).to(device)
        y = torch.randint(0, self.sizes[-1], (batch_size,), generator=generator).to(device)
        return x, y

    def flops_per_sample(self) -> float:
        return 6.0 * sum(self.sizes[i] * self.sizes[i + 1] for i in range(self.num_layers))


def mnist_mlp(**kw) -> MLP:
    """784-100-10 (BASELINE.json config 1)."""
    return MLP([784, 100, 10], **kw)


def wide_mlp(**kw) -> MLP:
    """784-4096x4-10 (BASELINE.json config 2): 53,600,266 parameters."""
    return MLP([784, 4096, 4096, 4096, 4096, 10], **kw)

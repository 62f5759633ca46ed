"""Model protocol shared by every scheme engine.

A model is a list of named parameters (TF-style names, so checkpoints are keyed
the way the reference's would be) plus ``forward_backward``: given the current
parameter tensors and a batch, return the loss and one gradient per parameter.
The reference's only "model" is the two-vector toy (Non-Distributed_Setup.py:12-18);
the MLP / GPT-2 / BERT / ResNet families come from BASELINE.json's configs.
"""
from __future__ import annotations

import dataclasses
from typing import Callable, Dict, List, Optional, Tuple

import torch

__all__ = ["ParamSpec", "Model"]


@dataclasses.dataclass
class ParamSpec:
    name: str
    shape: Tuple[int, ...]
    init: Callable[[Tuple[int, ...], Optional[torch.Generator]], torch.Tensor]
    #: the same initialiser as a tag the device-side kernel understands (``ops/init.py``):
    #: ("normal", std) | ("trunc_normal", std) | ("uniform", bound) | ("constant", value)
    dist: Optional[Tuple[str, float]] = None

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


class Model:
    """Reference (plain PyTorch, any device) implementation; native sm_100a
    versions subclass and override ``forward_backward``."""

    name = "model"

    def param_specs(self) -> List[ParamSpec]:
        raise NotImplementedError

    def init_params(self, seed: int = 0, device="cpu") -> Dict[str, torch.Tensor]:
        gen = torch.Generator().manual_seed(seed)
        return {s.name: s.init(s.shape, gen).to(device) for s in self.param_specs()}

    def loss(self, params: Dict[str, torch.Tensor], batch) -> torch.Tensor:
        raise NotImplementedError

    def forward_backward(self, params: Dict[str, torch.Tensor], batch
                         ) -> Tuple[torch.Tensor, Dict[str, torch.Tensor]]:
        leaves = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
        loss = self.loss(leaves, batch)
        names = list(leaves)
        grads = torch.autograd.grad(loss, [leaves[n] for n in names], allow_unused=True)
        return loss.detach(), {n: (torch.zeros_like(leaves[n]) if g is None else g)
                               for n, g in zip(names, grads)}

    def synthetic_batch(self, batch_size: int, device="cpu",
                        generator: Optional[torch.Generator] = None):
        return None

    def num_params(self) -> int:
        return sum(s.numel for s in self.param_specs())

    def flops_per_sample(self) -> float:
        """fwd+bwd FLOPs per sample (6 x matmul params by default)."""
        return 6.0 * self.num_params()

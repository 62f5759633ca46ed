"""The reference's toy problem: ``a, b in R^2``, ``c = a + b``,
``loss = mean((c - 100)^2)`` (Non-Distributed_Setup.py:12-18, repeated in every
scheme script).  Gradient is closed-form: ``dL/da_i = dL/db_i = c_i - 100``.
The notebooks use ``loss = |a - 100|`` on a single variable
(Basics-Tutorial/Local-then-Global-Variables.ipynb:107) -- ``ToyAbs``.
"""
from __future__ import annotations

import torch

from .base import Model, ParamSpec

__all__ = ["ToyModel", "ToyAbs"]


class ToyModel(Model):
    name = "toy"

    def __init__(self, init: str = "zeros", target: float = 100.0, dim: int = 2):
        self.init_kind, self.target, self.dim = init, float(target), dim

    def param_specs(self):
        if self.init_kind == "zeros":          # DOWNPOUR/DOWNPOUR.py:46-48
            f = lambda shape, gen: torch.zeros(shape)  # noqa: E731
        else:                                  # Non-Distributed_Setup.py:13-14
            def f(shape, gen):
                t = torch.empty(shape)
                torch.nn.init.trunc_normal_(t, 0.0, 1.0, -2.0, 2.0, generator=gen)
                return t
        dist = ("constant", 0.0) if self.init_kind == "zeros" else ("trunc_normal", 1.0)
        return [ParamSpec("Variable", (self.dim,), f, dist),
                ParamSpec("Variable_1", (self.dim,), f, dist)]

    def output(self, params):
        return params["Variable"] + params["Variable_1"]

    def loss(self, params, batch=None):
        return torch.mean((self.output(params) - self.target) ** 2)

    def forward_backward(self, params, batch=None):
        c = self.output(params).float()
        e = c - self.target
        g = e * (2.0 / self.dim)
        return torch.mean(e * e), {"Variable": g.clone(), "Variable_1": g.clone()}


class ToyAbs(Model):
    name = "toy_abs"

    def __init__(self, target: float = 100.0):
        self.target = float(target)

    def param_specs(self):
        return [ParamSpec("a", (), lambda shape, gen: torch.zeros(shape))]

    def loss(self, params, batch=None):
        return torch.abs(params["a"] - self.target)

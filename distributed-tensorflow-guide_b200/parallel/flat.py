"""Flat parameter layout shared by workers and ps shards.

The GPU engines keep every model as ONE flat fp32 buffer (plus a bf16 mirror for
the tensor-core GEMMs).  Variables are grouped by the ps task that owns them
(``placement.shard_variables`` -- the ``replica_device_setter`` equivalent,
Synchronous-SGD/ssgd.py:38-40), so each ps shard is a contiguous range of the
flat buffer: a pull is one streaming kernel per shard, a push from a wgrad
epilogue is a base pointer plus an offset, and the fused reduce-scatter kernel
can alternatively cut the same buffer into equal byte ranges.

Every variable starts on a 64-element boundary (128 B in bf16, 256 B in fp32) so
that TMA descriptors and 16 B vector accesses are always legal.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Sequence, Tuple

import torch

from ..placement import shard_variables

__all__ = ["FlatEntry", "FlatLayout"]

ALIGN = 64


@dataclasses.dataclass
class FlatEntry:
    name: str
    shape: Tuple[int, ...]          # logical (checkpoint) shape
    stored_shape: Tuple[int, ...]   # padded shape actually laid out
    offset: int                     # elements from the start of the flat buffer
    shard: int

    @property
    def numel(self) -> int:
        n = 1
        for s in self.stored_shape:
            n *= s
        return n


class FlatLayout:
    def __init__(self, specs: Sequence[Tuple[str, Tuple[int, ...], Tuple[int, ...]]],
                 num_shards: int = 1, strategy: str = "round_robin"):
        """``specs``: ``(name, logical_shape, stored_shape)`` in creation order."""
        sizes = []
        for name, shape, stored in specs:
            n = 1
            for s in stored:
                n *= s
            sizes.append((name, n * 4))
        self.num_shards = max(1, num_shards)
        self.assignment = shard_variables(sizes, self.num_shards, strategy)
        self.entries: Dict[str, FlatEntry] = {}
        self.order: List[str] = []
        self.shard_ranges: List[Tuple[int, int]] = []
        off = 0
        for s in range(self.num_shards):
            begin = off
            for name, shape, stored in specs:
                if self.assignment[name] != s:
                    continue
                e = FlatEntry(name, tuple(shape), tuple(stored), off, s)
                self.entries[name] = e
                self.order.append(name)
                off += -(-e.numel // ALIGN) * ALIGN
            self.shard_ranges.append((begin, off))
        self.total = off

    def view(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        e = self.entries[name]
        return flat[e.offset:e.offset + e.numel].view(e.stored_shape)

    def logical(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        e = self.entries[name]
        v = self.view(flat, name)
        if e.stored_shape != e.shape:
            v = v[tuple(slice(0, s) for s in e.shape)]
        return v

    def pack(self, flat: torch.Tensor, values: Dict[str, torch.Tensor]) -> None:
        """Write logical tensors into the flat buffer (padding stays zero)."""
        for name, val in values.items():
            if name in self.entries:
                self.logical(flat, name).copy_(val.to(flat.dtype))

    def unpack(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {n: self.logical(flat, n) for n in self.order}

    def shard_of(self, name: str) -> int:
        return self.entries[name].shard

    def offset_in_shard(self, name: str) -> int:
        e = self.entries[name]
        return e.offset - self.shard_ranges[e.shard][0]

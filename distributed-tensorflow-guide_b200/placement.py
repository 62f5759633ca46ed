"""``replica_device_setter`` and the parameter sharder.

Reference: ``tf.train.replica_device_setter(ps_tasks, worker_device)`` places
variable-creating ops on ``/job:ps/task:(i mod ps_tasks)`` round-robin in
creation order and everything else on ``worker_device``
(Synchronous-SGD/ssgd.py:38-40, DOWNPOUR/DOWNPOUR.py:86-87).  The reference
only ever uses one ps, and ``create_global_variables`` even hard-pins to
``/job:ps/task:0`` (DOWNPOUR/DOWNPOUR.py:202), so sharding is degenerate there.

On one 8xB200 host the ps tasks are HBM shards on (up to) 8 GPUs, so real
sharding matters: besides TF's round-robin this module offers a greedy
size-balanced strategy (largest variable first onto the lightest shard) and a
flat *byte-range* sharder used by the fused reduce-scatter/update/all-gather
kernel, where every GPU owns an equal, 16-byte-aligned slice of one flat
parameter buffer.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence, Tuple

from .cluster import ClusterSpec
from .framework import OpDescriptor

__all__ = ["replica_device_setter", "RoundRobinStrategy", "GreedyLoadBalancingStrategy",
           "shard_variables", "flat_shards", "byte_size_load_fn"]

_VARIABLE_OPS = ("Variable", "VariableV2", "VarHandleOp", "AutoReloadVariable",
                 "MutableHashTable", "MutableHashTableOfTensors")


class RoundRobinStrategy:
    """TF's default: task ``i mod num_tasks`` in creation order."""

    def __init__(self, num_tasks: int):
        self._n = max(1, int(num_tasks))
        self._next = 0

    def __call__(self, opdesc: OpDescriptor = None, nbytes: int = 0) -> int:
        t = self._next
        self._next = (self._next + 1) % self._n
        return t


class GreedyLoadBalancingStrategy:
    """Place each variable on the currently lightest ps task (by bytes)."""

    def __init__(self, num_tasks: int, load_fn: Optional[Callable] = None):
        self._loads = [0] * max(1, int(num_tasks))
        self._load_fn = load_fn

    def __call__(self, opdesc: OpDescriptor = None, nbytes: int = 1) -> int:
        if self._load_fn is not None and opdesc is not None:
            nbytes = self._load_fn(opdesc)
        t = min(range(len(self._loads)), key=lambda i: (self._loads[i], i))
        self._loads[t] += max(1, int(nbytes))
        return t

    @property
    def loads(self) -> List[int]:
        return list(self._loads)


def byte_size_load_fn(opdesc: OpDescriptor) -> int:
    return int(getattr(opdesc, "nbytes", 1))


def replica_device_setter(ps_tasks: int = 0, ps_device: str = "/job:ps",
                          worker_device: str = "/job:worker", merge_devices: bool = True,
                          cluster: Optional[ClusterSpec] = None,
                          ps_ops: Optional[Sequence[str]] = None,
                          ps_strategy: Optional[Callable] = None) -> Optional[Callable]:
    """Device function: variables -> ps tasks, the rest -> ``worker_device``."""
    if cluster is not None:
        cs = cluster if isinstance(cluster, ClusterSpec) else ClusterSpec(cluster)
        ps_tasks = cs.num_tasks("ps") if "ps" in cs.jobs else 0
    if ps_tasks == 0:
        return None
    ops = tuple(ps_ops) if ps_ops is not None else _VARIABLE_OPS
    strategy = ps_strategy or RoundRobinStrategy(ps_tasks)

    def chooser(opdesc: OpDescriptor) -> str:
        if opdesc.type in ops:
            return "%s/task:%d" % (ps_device, strategy(opdesc))
        return worker_device

    return chooser


# --------------------------------------------------------------------------- #
# Whole-model sharding (used by the GPU engines)
# --------------------------------------------------------------------------- #
def shard_variables(sizes: Sequence[Tuple[str, int]], num_shards: int,
                    strategy: str = "round_robin") -> Dict[str, int]:
    """Map variable name -> ps task.

    ``round_robin`` reproduces ``replica_device_setter`` (creation order);
    ``balanced`` sorts by size (largest first) and fills the lightest shard.
    """
    num_shards = max(1, int(num_shards))
    if strategy == "round_robin":
        return {name: i % num_shards for i, (name, _) in enumerate(sizes)}
    if strategy == "balanced":
        strat = GreedyLoadBalancingStrategy(num_shards)
        out = {}
        for name, nbytes in sorted(sizes, key=lambda kv: (-kv[1], kv[0])):
            out[name] = strat(None, nbytes)
        return out
    raise ValueError("unknown sharding strategy %r" % (strategy,))


def flat_shards(total_elems: int, num_shards: int, align_elems: int = 4) -> List[Tuple[int, int]]:
    """Equal ``[begin, end)`` element ranges of a flat buffer, each starting on
    an ``align_elems`` boundary (4 fp32 = 16 B, the vector width of the push /
    reduce kernels).  The last shard absorbs the remainder."""
    num_shards = max(1, int(num_shards))
    per = -(-total_elems // num_shards)
    per = -(-per // align_elems) * align_elems
    out = []
    for s in range(num_shards):
        b = min(s * per, total_elems)
        e = min(b + per, total_elems)
        out.append((b, e))
    return out

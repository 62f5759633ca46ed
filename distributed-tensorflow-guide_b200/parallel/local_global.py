"""Local <-> global variable helpers.

These are the only reusable functions of the reference
(DOWNPOUR/DOWNPOUR.py:149-233, repeated in DOWNPOUR-Easy/DOWNPOUR.py:142-202 and
ADAG/ADAG.py:138-190; prototyped in
Basics-Tutorial/Local-then-Global-Variables.ipynb:103-134):

* every worker owns *local* replicas of the model variables
  (``LOCAL_VARIABLES``), trains on them for a window of steps,
* a *global* copy named ``'g/' + local.op.name`` lives on the parameter server,
* ``assign_local_to_global`` seeds the ps from the chief, ``assign_global_to_local``
  pulls.

Difference by design: the reference pins every global copy to ``/job:ps/task:0``
even inside a ``replica_device_setter`` scope (DOWNPOUR/DOWNPOUR.py:202), which
defeats sharding.  Here the enclosing device scope decides (so the setter
shards round-robin); pass ``ps_device='/job:ps/task:0'`` for the literal
behaviour.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

from .. import framework as fw
from ..framework import GraphKeys, Op, Variable

__all__ = ["create_global_variables", "assign_global_to_local", "assign_local_to_global",
           "add_global_variables_to_local_collection", "clear_global_collection",
           "get_variable_by_name", "get_global_variable_by_name"]


def create_global_variables(local_optimizer_vars: Sequence[Variable] = (),
                            ps_device: Optional[str] = None
                            ) -> Tuple[Dict[Variable, Variable], Dict[Variable, Variable]]:
    """One global variable ``g/<name>`` per local variable, skipping the local
    optimizer's slots.  Returns ``(local_to_global, global_to_local)``."""
    local_to_global: Dict[Variable, Variable] = {}
    global_to_local: Dict[Variable, Variable] = {}
    skip = set(id(v) for v in local_optimizer_vars)

    def build():
        for v in fw.local_variables():
            if id(v) in skip:
                continue
            v_g = fw.get_variable("g/" + v.op.name, shape=v.shape, dtype=v.dtype,
                                  trainable=True,
                                  collections=[GraphKeys.GLOBAL_VARIABLES,
                                               GraphKeys.TRAINABLE_VARIABLES])
            local_to_global[v] = v_g
            global_to_local[v_g] = v

    if ps_device is not None:
        with fw.device(ps_device):
            build()
    else:
        build()
    return local_to_global, global_to_local


def assign_global_to_local(global_to_local: Dict[Variable, Variable]) -> Op:
    """PULL: ``local <- global`` for every pair (one batched read per ps task)."""
    pairs = list(global_to_local.items())

    def run(ctx):
        by_host = {}
        for g, l in pairs:
            by_host.setdefault((g._device_spec.job, g._device_spec.task), []).append((g, l))
        for group in by_host.values():
            vals = group[0][0]._host().read_many([g.storage_key for g, _ in group])
            for (_, l), val in zip(group, vals):
                l._assign_raw(val)
    return Op(run, "assign_global_to_local")


def assign_local_to_global(local_to_global: Dict[Variable, Variable]) -> Op:
    """Seed PUSH (chief): ``global <- local`` for every pair."""
    pairs = list(local_to_global.items())

    def run(ctx):
        by_host = {}
        for l, g in pairs:
            by_host.setdefault((g._device_spec.job, g._device_spec.task), []).append((l, g))
        for group in by_host.values():
            group[0][1]._host().assign_many([g.storage_key for _, g in group],
                                            [l.read_value() for l, _ in group])
    return Op(run, "assign_local_to_global")


def add_global_variables_to_local_collection() -> List[Variable]:
    """Move-what-the-local-optimizer-created trick, step 1
    (DOWNPOUR/DOWNPOUR.py:216-225)."""
    r = []
    for var in fw.get_collection(GraphKeys.GLOBAL_VARIABLES):
        fw.add_to_collection(GraphKeys.LOCAL_VARIABLES, var)
        if GraphKeys.LOCAL_VARIABLES not in var.collections:
            var.collections.append(GraphKeys.LOCAL_VARIABLES)
        r.append(var)
    return r


def clear_global_collection() -> None:
    """Step 2: empty ``GLOBAL_VARIABLES`` (DOWNPOUR/DOWNPOUR.py:228-233).  The
    variables that were in it become per-worker: they are re-homed onto this
    task so that two workers' local slots never alias on the ps."""
    from .. import runtime
    srv = runtime.current_server()
    ref = fw.get_collection_ref(GraphKeys.GLOBAL_VARIABLES)
    for var in list(ref):
        if GraphKeys.GLOBAL_VARIABLES in var.collections:
            var.collections.remove(GraphKeys.GLOBAL_VARIABLES)
        if GraphKeys.LOCAL_VARIABLES in var.collections:
            var._device_spec.job, var._device_spec.task = srv.job_name, srv.task_index
    del ref[:]


def get_variable_by_name(name: str) -> Variable:
    return [v for v in fw.get_collection("variables") if v.name == name][0]


def get_global_variable_by_name(name: str) -> Variable:
    return [v for v in fw.global_variables() if v.name == name][0]

"""Parameter stores: where the *global* (ps-resident) copy of a model lives and
how workers push to / pull from it.

``HostStore`` is the CPU / control-plane implementation: global variables are
``Variable`` objects named ``g/<name>`` (DOWNPOUR/DOWNPOUR.py:205) -- or the
plain name for schemes that keep a single shared copy (Hogwild/Hogwild.py:37-38,
Synchronous-SGD/ssgd.py:42-43) -- sharded over the cluster's ps tasks by a
``replica_device_setter``-equivalent (placement.shard_variables), and every
update executes on the owning task (rpc.VariableHost).

``parallel/peer_store.py`` implements the same interface on B200 with the ps
shards in NVLink-mapped HBM and pushes/pulls done by sm_100a kernels.

Interface used by the scheme engines (``parallel/engines.py``):

``init(is_chief, values)``, ``pull()``, ``push(grads, rule, lr, scale)``,
``bump_step()``, ``read_step()``, ``setup_sync(...)``, ``sync_step(grads,
scale)``, ``start_chief_services()``, ``stop_chief_services()``,
``global_variables()``, ``mark_done()``.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch

from .. import framework as fw
from .. import runtime
from ..framework import GraphKeys, Variable
from ..models.base import Model
from ..placement import shard_variables
from ..train.optimizers import (AdagradOptimizer, GradientDescentOptimizer, Optimizer,
                                SyncReplicasOptimizer)

__all__ = ["HostStore"]


class HostStore:
    def __init__(self, model: Model, num_ps: int = 1, prefix: str = "g/",
                 sharding: str = "round_robin", ps_job: Optional[str] = "ps",
                 create_global_step: bool = True):
        self.model = model
        self.specs = model.param_specs()
        self.prefix = prefix
        self.num_ps = max(1, num_ps)
        sizes = [(s.name, s.numel * 4) for s in self.specs]
        self.assignment = shard_variables(sizes, self.num_ps, sharding)
        self.vars: Dict[str, Variable] = {}
        srv = runtime.current_server()
        has_ps = ps_job is not None and getattr(srv, "cluster", None) is not None \
            and ps_job in srv.cluster.jobs
        g = fw.get_default_graph()
        saved, g._name_scope = g._name_scope, []
        try:
            for s in self.specs:
                dev = "/job:%s/task:%d" % (ps_job, self.assignment[s.name]) if has_ps else None
                with fw.device(None), fw.device(dev):
                    self.vars[s.name] = Variable(
                        torch.zeros(s.shape), name=prefix + s.name, trainable=True,
                        collections=[GraphKeys.GLOBAL_VARIABLES, GraphKeys.TRAINABLE_VARIABLES])
            self.global_step: Optional[Variable] = None
            if create_global_step:
                dev = "/job:%s/task:0" % ps_job if has_ps else None
                with fw.device(None), fw.device(dev):
                    self.global_step = fw.get_global_step() or Variable(
                        torch.zeros((), dtype=torch.int64), trainable=False, name="global_step")
        finally:
            g._name_scope = saved
        self._opts: Dict[Tuple[str, float], Optimizer] = {}
        self._sync: Optional[SyncReplicasOptimizer] = None
        self._sync_grads: Dict[str, torch.Tensor] = {}
        self._sync_train = None
        self._sync_ctx_session = None
        self.bytes_pushed = 0
        self.bytes_pulled = 0

    # ---------------------------------------------------------------- basics
    @property
    def param_names(self) -> List[str]:
        return [s.name for s in self.specs]

    def global_variables(self) -> List[Variable]:
        out = list(self.vars.values())
        if self.global_step is not None:
            out.append(self.global_step)
        return out

    def _by_host(self, names: Sequence[str]):
        groups: Dict[Tuple, List[str]] = {}
        for n in names:
            v = self.vars[n]
            groups.setdefault((v._device_spec.job, v._device_spec.task), []).append(n)
        return groups

    def seed(self, values: Dict[str, torch.Tensor]) -> None:
        """Chief: ps <- chief's initial values (K6, DOWNPOUR/DOWNPOUR.py:130)."""
        for names in self._by_host(values.keys()).values():
            self.vars[names[0]]._host().assign_many(
                [self.vars[n].storage_key for n in names],
                [values[n].detach().float().cpu() for n in names])

    def pull(self, device="cpu") -> Dict[str, torch.Tensor]:
        """PULL (K5): one batched read per ps task."""
        out = {}
        for names in self._by_host(self.param_names).values():
            vals = self.vars[names[0]]._host().read_many(
                [self.vars[n].storage_key for n in names])
            for n, v in zip(names, vals):
                out[n] = v.to(device)
                self.bytes_pulled += v.numel() * v.element_size()
        return out

    def _optimizer(self, rule: str, lr: float) -> Optimizer:
        key = (rule, float(lr))
        opt = self._opts.get(key)
        if opt is None:
            if rule == "sgd":
                opt = GradientDescentOptimizer(lr)
            elif rule == "adagrad":
                opt = AdagradOptimizer(lr)
                # share slots between lr values: one accumulator per variable
                for o in self._opts.values():
                    if isinstance(o, AdagradOptimizer):
                        opt._slots = o._slots
                        break
                opt._create_slots(list(self.vars.values()))
            else:
                raise ValueError("unknown update rule %r" % rule)
            self._opts[key] = opt
        return opt

    def declare_rule(self, rule: str, lr: float) -> None:
        """Create slot variables up front so the chief's init op covers them."""
        self._optimizer(rule, lr)

    def push(self, grads: Dict[str, torch.Tensor], rule: str = "sgd", lr: float = 1e-4,
             scale: float = 1.0) -> None:
        """PUSH (K3/K4): apply ``scale*grad`` on the ps with the given rule;
        unlocked, one request per ps task."""
        opt = self._optimizer(rule, lr)
        for names in self._by_host(grads.keys()).values():
            vs = [self.vars[n] for n in names]
            gs = [(grads[n].detach().float().cpu() * scale) if scale != 1.0
                  else grads[n].detach().float().cpu() for n in names]
            opt._apply_on_host(vs[0]._host(), vs, gs)
            self.bytes_pushed += sum(g.numel() * 4 for g in gs)

    def bump_step(self, n: int = 1) -> int:
        return self.global_step._host().counter_add(self.global_step.storage_key, n)

    def read_step(self) -> int:
        return int(self.global_step.read_value().item())

    def mark_done(self, who: str) -> None:
        srv = runtime.current_server()
        cluster = getattr(srv, "cluster", None)
        if cluster is None or "ps" not in cluster.jobs:
            return
        for t in cluster.task_indices("ps"):
            try:
                srv.host_for("ps", t).mark_done(who)
            except Exception:
                pass

    # ------------------------------------------------------------------ sync
    def setup_sync(self, rule: str, lr: float, replicas_to_aggregate: int,
                   total_num_replicas: int) -> SyncReplicasOptimizer:
        """SyncReplicas plumbing over this store's variables; gradients are fed
        from ``sync_step``."""
        base = self._optimizer(rule, lr)
        self._sync = SyncReplicasOptimizer(base, replicas_to_aggregate, total_num_replicas)
        gv = [(fw.tensor(lambda n=n: self._sync_grads[n], "sync_grad/" + n), self.vars[n])
              for n in self.param_names]
        self._sync_train = self._sync.apply_gradients(gv, global_step=self.global_step)
        return self._sync

    def sync_step(self, grads: Dict[str, torch.Tensor], scale: float = 1.0, session=None) -> int:
        """Push to the accumulators (stale ones are dropped), block for a token,
        return the new global step carried by the token."""
        self._sync_grads = {n: (g.detach().float().cpu() * scale) for n, g in grads.items()}
        from ..framework import RunContext, evaluate
        evaluate(self._sync_train, RunContext(session))
        return int(self._sync._local_step.read_value().item())

    @property
    def sync_optimizer(self) -> Optional[SyncReplicasOptimizer]:
        return self._sync

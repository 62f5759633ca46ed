"""Scheme engines: the training algorithms of the reference as first-class
objects (SURVEY Appendix B gives them in one place).

====================  =========================================================
``single``            Non-Distributed_Setup.py:12-30 -- plain SGD.
``hogwild``           Hogwild/Hogwild.py:36-57 -- every worker: pull, grad,
                      unlocked ``w_ps -= lr*g``.  No barrier of any kind.
``downpour``          DOWNPOUR/DOWNPOUR.py:63-102 -- T local Adagrad steps, the
                      *sum* of the window's grads pushed with Adagrad on the ps,
                      then pull.
``downpour_easy``     DOWNPOUR-Easy/DOWNPOUR.py:53-94 -- same with local SGD at a
                      per-worker rate (``lr*task_index`` in the reference).
``adag``              ADAG/ADAG.py:69-90 -- local SGD, window *mean*, SGD on ps.
``ssgd``              Synchronous-SGD/ssgd.py:51-62 -- N-of-M mean, stale drop.
``ssgd_lr``           Synchronous-SGD-different-learning-rates/ssgd.py:51-65 --
                      worker k pre-scales by ``(lr_k/base_lr)*N``; ps applies the
                      mean with ``base_lr`` => ``w -= sum_k lr_k g_k``.
``sdag``              SDAG/README.md:3 intent: T local steps, window mean, then
                      synchronous aggregation (the script itself is WIP and has
                      no local updates, SDAG/dist_cpu_sing_mach_sync.py:60-68).
``multigpu_async``    Multiple-GPUs-Single-Machine/dist_mult_gpu_sing_mach.py --
                      Hogwild with one GPU per worker.
``tower_avg``         BASELINE.json north star: per-step mean of per-GPU grads,
                      one update, replicas stay identical.
====================  =========================================================

Deliberate deltas from the scripts: the window does *true* local
steps (the TF graphs evaluate the forward once per ``run``, #2); ``ssgd_lr``
uses ``N`` instead of the literal ``2.0`` (#9); tokens are seeded once (#8).

The engines are backend-agnostic: they talk to a *store* (``HostStore`` on
CPU, ``PeerStore`` on B200) and a *model* (plain PyTorch or the native sm_100a
implementation), so the same scheme code drives the toy problem on localhost
and GPT-2 on eight GPUs.
"""
from __future__ import annotations

import dataclasses
import time
from typing import Callable, Dict, List, Optional, Sequence, Union

import torch

from .. import framework as fw
from .. import runtime
from ..framework import Op, Variable
from ..models.base import Model
from ..session.hooks import SessionRunHook, StopAtStepHook
from ..session.monitored import MonitoredTrainingSession, Scaffold
from ..session.checkpoint import Saver

__all__ = ["SchemeConfig", "WorkerEngine", "SCHEMES", "LocalOptimizerState"]

SCHEMES = ("single", "hogwild", "downpour", "downpour_easy", "adag", "ssgd", "ssgd_lr",
           "sdag", "multigpu_async", "tower_avg",
           # the reference's TODO list (README.md:40-42), built here:
           "dynsgd", "aeasgd", "aeamsgd")


@dataclasses.dataclass
class SchemeConfig:
    scheme: str = "hogwild"
    lr: float = 1e-4                      # ps-side rate (base rate for ssgd_lr)
    local_lr: Union[None, float, Sequence[float]] = None   # per-worker list allowed
    update_window: int = 1                # T
    replicas_to_aggregate: Optional[int] = None
    total_num_replicas: Optional[int] = None
    worker_lrs: Optional[Sequence[float]] = None            # ssgd_lr
    last_step: Optional[int] = None       # StopAtStepHook on global_step
    max_iters: Optional[int] = None       # plain loop bound (Hogwild's 1000)
    checkpoint_dir: Optional[str] = None
    save_checkpoint_secs: Optional[float] = None
    ps_sharding: str = "round_robin"
    stop_grace_period_secs: float = 10.0
    seed: int = 0
    num_tokens: int = -1                  # initial sync tokens (-1 = replicas_to_aggregate, TF default)
    elastic_alpha: Optional[float] = None  # AEASGD moving rate (default 0.9 / n_workers)
    momentum: float = 0.9                 # AEAMSGD Nesterov momentum (delta)

    def resolved(self, n_workers: int) -> "SchemeConfig":
        c = dataclasses.replace(self)
        s = c.scheme
        if s not in SCHEMES:
            raise ValueError("unknown scheme %r (have: %s)" % (s, ", ".join(SCHEMES)))
        if s in ("downpour", "downpour_easy", "adag") and c.update_window == 1:
            c.update_window = 3           # DOWNPOUR.py:63, ADAG.py:69
        if s == "sdag" and c.update_window == 1:
            c.update_window = 5           # SDAG/dist_cpu_sing_mach_sync.py:60
        if s in ("aeasgd", "aeamsgd", "dynsgd") and c.update_window == 1:
            c.update_window = 4           # communication period tau
        if s in ("aeasgd", "aeamsgd") and c.elastic_alpha is None:
            c.elastic_alpha = 0.9 / max(1, n_workers)     # alpha = beta / p, beta = 0.9
        if s in ("ssgd", "ssgd_lr", "sdag", "tower_avg"):
            c.total_num_replicas = c.total_num_replicas or n_workers
            c.replicas_to_aggregate = c.replicas_to_aggregate or c.total_num_replicas
        if s == "ssgd_lr":
            if c.worker_lrs is None:      # ssgd.py:59-62: task 0 -> 0.1, others 1e-4
                c.worker_lrs = [0.1] + [1e-4] * (n_workers - 1)
            c.lr = 1.0                    # base_lr, ssgd.py:51
        return c

    def local_lr_for(self, task_index: int) -> float:
        if self.local_lr is None:
            if self.scheme == "downpour_easy":   # DOWNPOUR-Easy/DOWNPOUR.py:53
                return self.lr * task_index
            return self.lr
        if isinstance(self.local_lr, (int, float)):
            return float(self.local_lr)
        return float(self.local_lr[task_index])


class LocalOptimizerState:
    """Worker-local optimizer + window accumulator (K7 + K8).

    Plain-torch implementation; the native path (``ops.fused_local_step``) does
    the same in one multi-tensor kernel over flat buffers.
    """

    def __init__(self, rule: str, lr: float, params: Dict[str, torch.Tensor],
                 initial_accumulator_value: float = 0.1):
        self.rule, self.lr = rule, lr
        self.slots = {n: torch.full_like(p, initial_accumulator_value, dtype=torch.float32)
                      for n, p in params.items()} if rule == "adagrad" else {}
        self.window: Dict[str, torch.Tensor] = {}
        self.local_step = 0

    def reset_window(self, params):
        self.window = {n: torch.zeros_like(p, dtype=torch.float32) for n, p in params.items()}

    def step(self, params: Dict[str, torch.Tensor], grads: Dict[str, torch.Tensor]) -> None:
        for n, g in grads.items():
            g = g.float()
            self.window[n] += g
            p = params[n]
            if self.rule == "adagrad":
                acc = self.slots[n]
                acc.addcmul_(g, g)
                p.sub_((self.lr * g / acc.sqrt()).to(p.dtype))
            else:
                p.sub_((self.lr * g).to(p.dtype))
        self.local_step += 1


class WorkerEngine:
    """One worker's program for any scheme.

    Typical use (mirrors the scripts' flow, e.g. DOWNPOUR/DOWNPOUR.py:116-146)::

        eng = WorkerEngine(model, cfg, task_index=k, n_workers=2, store=store)
        with eng.session(master=server.target) as sess:
            while not sess.should_stop():
                sess.run(eng.train_op)
    """

    def __init__(self, model: Model, cfg: SchemeConfig, task_index: int = 0, n_workers: int = 1,
                 store=None, batch_fn: Optional[Callable[[], object]] = None, device="cpu",
                 batch_size: int = 0):
        self.model = model
        self.cfg = cfg.resolved(n_workers)
        self.task_index, self.n_workers = task_index, n_workers
        self.is_chief = task_index == 0
        self.device = device
        self.batch_fn = batch_fn or (lambda: model.synthetic_batch(batch_size, device)
                                     if batch_size else None)
        if store is None:
            from .store import HostStore
            srv = runtime.current_server()
            cluster = getattr(srv, "cluster", None)
            n_ps = cluster.num_tasks("ps") if cluster is not None and "ps" in cluster.jobs else 1
            prefix = "g/" if self.cfg.scheme in ("downpour", "downpour_easy", "adag") else ""
            store = HostStore(model, n_ps, prefix=prefix, sharding=self.cfg.ps_sharding)
        self.store = store
        self.local: Dict[str, torch.Tensor] = {}
        self.last_loss: Optional[torch.Tensor] = None
        self.iters = 0
        self._lopt: Optional[LocalOptimizerState] = None
        self._elastic_x = None
        self._elastic_v = None
        self.last_staleness = 1
        from ..utils.trace import PhaseTracer
        self.tracer = PhaseTracer.from_env(rank=task_index)     # host-clock phases (DTG_TRACE=1)
        self._build()

    # ------------------------------------------------------------------ build
    def _build(self):
        c, s = self.cfg, self.cfg.scheme
        if s in ("downpour", "downpour_easy"):
            self.store.declare_rule("adagrad", c.lr)          # ps slots g/<v>/Adagrad
        if s in ("ssgd", "ssgd_lr", "sdag", "tower_avg"):
            self._sync = self.store.setup_sync("sgd", c.lr, c.replicas_to_aggregate,
                                               c.total_num_replicas)
        else:
            self._sync = None
        # local step counter, like the scripts' 'local_step' in 'local_non_trainable'
        with fw.device(None):
            self.local_step_var = Variable(torch.zeros((), dtype=torch.int64), trainable=False,
                                           name="local_step", collections=["local_non_trainable"])
        step_fn = {
            "single": self._step_hogwild, "hogwild": self._step_hogwild,
            "multigpu_async": self._step_hogwild,
            "downpour": self._step_window_async, "downpour_easy": self._step_window_async,
            "adag": self._step_window_async,
            "ssgd": self._step_sync, "ssgd_lr": self._step_sync, "tower_avg": self._step_sync,
            "sdag": self._step_sync_window,
            "dynsgd": self._step_dynsgd, "aeasgd": self._step_elastic,
            "aeamsgd": self._step_elastic,
        }[s]
        self._session = None
        self.train_op: Op = Op(lambda ctx: step_fn(getattr(ctx, "session", None)), "train/" + s)
        self.loss = fw.tensor(lambda: self.last_loss, "loss")
        self.global_step = self.store.global_step
        # --- init stages ----------------------------------------------------
        init_vals = self.model.init_params(seed=c.seed)

        def chief_init(ctx):
            # chief initialises GLOBAL variables: model values, slots, step
            self.store.seed(init_vals)
            for v in fw.global_variables():
                if v not in self.store.vars.values() and not v.is_initialized():
                    v._assign_raw(v.initial_value())

        def local_init(ctx):
            self.local_step_var._assign_raw(0)
            for v in fw.local_variables():
                v._assign_raw(v.initial_value())

        self.init_op = Op(chief_init, "init")
        local_ops = [Op(local_init, "init_local")]
        if self._sync is not None:
            local_ops.append(self._sync.chief_init_op if self.is_chief
                             else self._sync.local_step_init_op)
        self.scaffold = Scaffold(init_op=self.init_op, local_init_op=local_ops,
                                 saver=Saver(var_list=None))
        self.hooks: List[SessionRunHook] = []
        if self._sync is not None:
            self.hooks.append(self._sync.make_session_run_hook(self.is_chief, c.num_tokens))
        if c.last_step is not None:
            self.hooks.append(StopAtStepHook(last_step=c.last_step))

    def session(self, master: str = "", config=None, extra_hooks: Sequence[SessionRunHook] = ()):
        c = self.cfg
        sess = MonitoredTrainingSession(
            master=master, is_chief=self.is_chief, config=config, scaffold=self.scaffold,
            hooks=list(self.hooks) + list(extra_hooks), checkpoint_dir=c.checkpoint_dir,
            save_checkpoint_secs=c.save_checkpoint_secs if c.checkpoint_dir else None,
            stop_grace_period_secs=c.stop_grace_period_secs)
        self._session = sess
        # every worker starts from the ps state ("grab_global_init", DOWNPOUR.py:135)
        self.local = self.store.pull(self.device)
        return sess

    # ------------------------------------------------------------- step rules
    def _fb(self):
        loss, grads = self.model.forward_backward(self.local, self.batch_fn())
        self.last_loss = loss
        self.iters += 1
        return grads

    def _step_hogwild(self, session=None):
        """pull -> grad -> unlocked SGD on ps (K5, K1/K2, K3)."""
        tr = self.tracer
        with tr.range("pull"):
            self.local = self.store.pull(self.device)
        with tr.range("fwd+bwd"):
            grads = self._fb()
        with tr.range("push"):
            self.store.push(grads, "sgd", self.cfg.lr)
        self.store.bump_step()
        self.local_step_var.assign_add(1)
        tr.step_done()

    def _window(self, rule: str, lr: float):
        T = self.cfg.update_window
        if self._lopt is None:
            self._lopt = LocalOptimizerState(rule, lr, self.local)
        self._lopt.reset_window(self.local)
        for _ in range(T):
            self._lopt.step(self.local, self._fb())
        self.local_step_var.assign_add(T)
        return self._lopt.window

    def _step_window_async(self, session=None):
        """One communication round of DOWNPOUR / DOWNPOUR-Easy / ADAG."""
        c, s = self.cfg, self.cfg.scheme
        local_rule = "adagrad" if s == "downpour" else "sgd"
        tr = self.tracer
        with tr.range("local_steps"):
            G = self._window(local_rule, c.local_lr_for(self.task_index))
        with tr.range("push"):
            if s == "adag":
                self.store.push(G, "sgd", c.lr, scale=1.0 / c.update_window)    # mean
            else:
                self.store.push(G, "adagrad", c.lr)                              # sum
        self.store.bump_step()
        with tr.range("pull"):
            self.local = self.store.pull(self.device)
        tr.step_done()

    def _step_dynsgd(self, session=None):
        """Dynamic SGD (staleness-aware async SGD; reference TODO, README.md:40).

        Pull and remember the ps version; run T local SGD steps accumulating the window sum;
        push ``lr * G / staleness`` where staleness = updates the ps absorbed since the pull
        + 1 -- a stale worker's contribution shrinks instead of being dropped."""
        c = self.cfg
        self.local = self.store.pull(self.device)
        seen = int(self.store.global_step.read_value())
        G = self._window("sgd", c.local_lr_for(self.task_index))
        staleness = max(1, int(self.store.global_step.read_value()) - seen + 1)
        self.last_staleness = staleness
        self.store.push(G, "sgd", c.lr, scale=1.0 / staleness)
        self.store.bump_step()

    def _step_elastic(self, session=None):
        """AEASGD / AEAMSGD (Zhang, Choromanska, LeCun 2015; reference TODO, README.md:41-42).

        Every tau local steps the worker and the centre variable on the ps pull each other:
        ``e = alpha (x_i - x~);  x_i -= e;  x~ += e``.  Between exchanges the worker runs plain
        SGD (AEASGD) or Nesterov momentum SGD evaluated at ``x + delta v`` (AEAMSGD)."""
        c = self.cfg
        if self._elastic_x is None:                       # first round: start from the centre
            self._elastic_x = {n: p.clone().float() for n, p in self.store.pull(self.device).items()}
            self._elastic_v = {n: torch.zeros_like(p) for n, p in self._elastic_x.items()}
        x, v = self._elastic_x, self._elastic_v
        centre = self.store.pull(self.device)
        e = {n: c.elastic_alpha * (x[n] - centre[n].float()) for n in x}
        for n in x:
            x[n] -= e[n]
        self.store.push(e, "sgd", 1.0, scale=-1.0)        # x~ += e  (unlocked, like every push)
        self.store.bump_step()
        lr = c.local_lr_for(self.task_index)
        for _ in range(c.update_window):
            if c.scheme == "aeamsgd":
                self.local = {n: x[n] + c.momentum * v[n] for n in x}
                g = self._fb()
                for n in x:
                    v[n].mul_(c.momentum).sub_(lr * g[n].float())
                    x[n] += v[n]
            else:
                self.local = x
                g = self._fb()
                for n in x:
                    x[n] -= lr * g[n].float()
        self.local = x
        self.local_step_var.assign_add(c.update_window)

    def _sync_scale(self) -> float:
        c = self.cfg
        if c.scheme == "ssgd_lr":
            return (c.worker_lrs[self.task_index] / c.lr) * c.replicas_to_aggregate
        return 1.0

    def _step_sync(self, session=None):
        """SSGD / per-worker-LR / tower averaging."""
        self.local = self.store.pull(self.device)
        grads = self._fb()
        self.store.sync_step(grads, self._sync_scale(), session=session)
        self.local_step_var.assign_add(1)

    def _step_sync_window(self, session=None):
        """SDAG: T local SGD steps, window mean, synchronous aggregation."""
        c = self.cfg
        self.local = self.store.pull(self.device)
        G = self._window("sgd", c.local_lr_for(self.task_index))
        self.store.sync_step(G, 1.0 / c.update_window, session=session)

    # ----------------------------------------------------------------- driver
    def output(self):
        """The scripts' fetched ``c``: model output on the *ps* parameters."""
        params = self.store.pull(self.device)
        if hasattr(self.model, "output"):
            return self.model.output(params)
        return None

    def run(self, master: str = "", log: Optional[Callable[[str], None]] = None,
            log_every: int = 0, pace_secs: float = 0.0, pre_sleep: float = 0.0,
            post_sleep: float = 0.0, extra_hooks: Sequence[SessionRunHook] = ()) -> dict:
        """Whole-worker program with the reference's structure: session
        bring-up, optional chief grace sleep, loop until ``should_stop`` or
        ``max_iters``, close, report done to the ps."""
        c = self.cfg
        t0 = time.time()
        sess = self.session(master, extra_hooks=extra_hooks)
        if pre_sleep and self.is_chief:
            time.sleep(pre_sleep)
        rounds = 0
        try:
            while not sess.should_stop():
                if c.max_iters is not None and rounds >= c.max_iters:
                    break
                sess.run(self.train_op)
                rounds += 1
                if log and log_every and rounds % log_every == 0:
                    out = self.output()
                    log("%s global step: %d worker: %d local step: %d" % (
                        out.tolist() if out is not None else float(self.last_loss),
                        self.store.read_step(), self.task_index,
                        int(self.local_step_var.read_value().item())))
                if pace_secs:
                    time.sleep(pace_secs)
        finally:
            if post_sleep:
                time.sleep(post_sleep)
            final_step = self.store.read_step()
            sess.close()
            self.store.mark_done("worker:%d" % self.task_index)
        return {"rounds": rounds, "iters": self.iters, "global_step": final_step,
                "seconds": time.time() - t0,
                "loss": None if self.last_loss is None else float(self.last_loss)}

"""Optimizers: SGD, Adagrad and the SyncReplicas wrapper.

Reference usage:

* ``GradientDescentOptimizer(lr).minimize(loss)`` -- Non-Distributed_Setup.py:20,
  Hogwild/Hogwild.py:44 (unlocked ``w -= lr*g`` on the ps).
* ``AdagradOptimizer(lr)`` locally and on the ps, ``compute_gradients(loss,
  var_list=tf.local_variables())`` / ``apply_gradients(zip(grads, vars),
  global_step=...)`` -- DOWNPOUR/DOWNPOUR.py:57,68-75,92-98.  Slot ``<var>/Adagrad``
  starts at 0.1 and is colocated with its variable.
* ``SyncReplicasOptimizer(opt, replicas_to_aggregate, total_num_replicas)`` with
  ``make_session_run_hook``, ``get_init_tokens_op``, ``chief_init_op``,
  ``local_step_init_op``, ``ready_for_local_init_op`` --
  Synchronous-SGD/ssgd.py:52-60,
  Synchronous-SGD-different-learning-rates/ssgd.py:51-75.

The update itself always executes on the task that owns the variable
(``VariableHost.apply_sgd`` / ``apply_adagrad``); on B200 those host methods are
replaced by the push kernels of ``ops`` writing into peer HBM (see
``parallel/engine.py``), the optimizer classes stay the user-facing surface.
"""
from __future__ import annotations

import threading
import time
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from .. import framework as fw
from ..framework import (GraphKeys, Node, Op, OutOfRangeError, RunContext, Tensor, Variable,
                         _GradContext, evaluate, get_global_step, get_or_create_global_step,
                         global_step_value)

__all__ = ["Optimizer", "GradientDescentOptimizer", "AdagradOptimizer",
           "SyncReplicasOptimizer", "OutOfRangeError", "get_global_step",
           "get_or_create_global_step", "global_step_value"]


class _Grads:
    """One ``compute_gradients`` call = one forward/backward per run."""

    def __init__(self, loss, var_list: Sequence[Variable]):
        self.loss, self.vars = loss, list(var_list)
        self._key = object()

    def values(self, ctx: RunContext) -> List[torch.Tensor]:
        k = id(self._key)
        if k in ctx.cache:
            return ctx.cache[k]
        gctx = _GradContext(ctx)
        if callable(self.loss) and not isinstance(self.loss, Node):
            leaves = [gctx.read(v) for v in self.vars]
            loss = self.loss(*leaves)
        else:
            loss = evaluate(self.loss, gctx)
            leaves = [gctx.read(v) for v in self.vars]
        grads = torch.autograd.grad(loss, leaves, allow_unused=True)
        out = [torch.zeros_like(l) if g is None else g.detach()
               for g, l in zip(grads, leaves)]
        ctx.cache[k] = out
        return out


class Optimizer:
    slot_name: Optional[str] = None

    def __init__(self, learning_rate: float, use_locking: bool = False, name: str = "Optimizer"):
        self._lr = float(learning_rate)
        self._use_locking = use_locking
        self._name = name
        self._slots: Dict[str, Variable] = {}

    # ------------------------------------------------------------------ grads
    def compute_gradients(self, loss, var_list: Optional[Sequence[Variable]] = None
                          ) -> List[Tuple[Tensor, Variable]]:
        """``loss`` is a lazy ``Tensor`` built from variables, or a callable
        ``loss(*var_values)``.  Returns ``[(grad, var), ...]`` lazily; all
        grads of one call share a single forward/backward, and each call reads
        the variables afresh when it is evaluated."""
        if var_list is None:
            var_list = fw.trainable_variables()
        var_list = list(var_list)
        if not var_list:
            raise ValueError("No variables to optimize.")
        g = _Grads(loss, var_list)
        out = []
        for i, v in enumerate(var_list):
            out.append((Tensor(lambda ctx, i=i: g.values(ctx)[i],
                               "gradients/%s_grad" % v.op.name), v))
        return out

    # ------------------------------------------------------------------ apply
    def _create_slots(self, var_list: Sequence[Variable]) -> None:
        pass

    def _apply_on_host(self, host, vars_: List[Variable], grads: List[torch.Tensor]) -> None:
        raise NotImplementedError

    def apply_gradients(self, grads_and_vars, global_step: Optional[Variable] = None,
                        name: Optional[str] = None) -> Op:
        gv = [(g, v) for g, v in grads_and_vars if g is not None]
        if not gv:
            raise ValueError("No gradients provided for any variable.")
        self._create_slots([v for _, v in gv])

        def run(ctx):
            by_host: Dict[Tuple, Tuple[list, list]] = {}
            for g, v in gv:
                key = (v._device_spec.job, v._device_spec.task)
                vs, gs = by_host.setdefault(key, ([], []))
                vs.append(v)
                gval = evaluate(g, ctx)
                gs.append(torch.as_tensor(gval).detach().to(torch.float32))
            for _, (vs, gs) in by_host.items():
                self._apply_on_host(vs[0]._host(), vs, gs)
            if global_step is not None:
                global_step._host().counter_add(global_step.storage_key, 1)
        return Op(run, name or self._name)

    def minimize(self, loss, global_step: Optional[Variable] = None,
                 var_list: Optional[Sequence[Variable]] = None, name: Optional[str] = None) -> Op:
        return self.apply_gradients(self.compute_gradients(loss, var_list),
                                    global_step=global_step, name=name)

    def get_slot(self, var: Variable, name: str) -> Optional[Variable]:
        return self._slots.get("%s/%s" % (var.op.name, name))

    def get_slot_names(self) -> List[str]:
        return [self.slot_name] if self.slot_name else []


class GradientDescentOptimizer(Optimizer):
    """``var -= lr * grad``; no slots (Non-Distributed_Setup.py:20)."""

    def __init__(self, learning_rate, use_locking=False, name="GradientDescent"):
        super().__init__(learning_rate, use_locking, name)

    def _apply_on_host(self, host, vars_, grads):
        host.apply_sgd([v.storage_key for v in vars_], grads, self._lr)


class AdagradOptimizer(Optimizer):
    """``accum += g*g; var -= lr*g/sqrt(accum)``; slot ``<var>/Adagrad`` = 0.1,
    colocated with the variable (DOWNPOUR/DOWNPOUR.py:57,92)."""

    slot_name = "Adagrad"

    def __init__(self, learning_rate, initial_accumulator_value=0.1, use_locking=False,
                 name="Adagrad"):
        super().__init__(learning_rate, use_locking, name)
        self._init_acc = float(initial_accumulator_value)

    def _create_slots(self, var_list):
        g = fw.get_default_graph()
        for v in var_list:
            key = "%s/%s" % (v.op.name, self._name)
            if key in self._slots:
                continue
            existing = g._variables_by_name.get(key)
            if existing is not None:
                self._slots[key] = existing
                continue
            saved_scope, g._name_scope = g._name_scope, []
            try:
                with fw.device(None), fw.device(v.device):
                    # [TF] slots are created in GLOBAL_VARIABLES -- which is why
                    # DOWNPOUR has to move the *local* optimizer's slots into the
                    # local collection by hand (DOWNPOUR/DOWNPOUR.py:80-84).
                    self._slots[key] = Variable(
                        torch.full(v.shape, self._init_acc, dtype=torch.float32),
                        trainable=False, name=key,
                        collections=[GraphKeys.GLOBAL_VARIABLES])
            finally:
                g._name_scope = saved_scope

    def _apply_on_host(self, host, vars_, grads):
        host.apply_adagrad([v.storage_key for v in vars_],
                           [self._slots["%s/%s" % (v.op.name, self._name)].storage_key
                            for v in vars_],
                           grads, self._lr, self._init_acc)


# --------------------------------------------------------------------------- #
# SyncReplicas
# --------------------------------------------------------------------------- #
class SyncReplicasOptimizer(Optimizer):
    """N-of-M synchronous aggregation with stale-gradient drop.

    Per variable there is a conditional accumulator on the variable's task;
    each worker's train op pushes its gradient tagged with its ``local_step``
    (dropped when older than the accumulator's step) and then blocks on a token
    from the shared queue; the chief's background thread takes the mean of the
    first ``replicas_to_aggregate`` gradients, applies the wrapped optimizer on
    the ps, bumps ``global_step`` and releases ``max(total, N)`` tokens carrying
    the new step (Synchronous-SGD/ssgd.py:51-62 and README.md:3 there).
    """

    TOKEN_QUEUE = "sync_token_q"

    def __init__(self, opt: Optimizer, replicas_to_aggregate: int,
                 total_num_replicas: Optional[int] = None, use_locking=False,
                 name="sync_replicas"):
        super().__init__(opt._lr, use_locking, name)
        self._opt = opt
        self._replicas_to_aggregate = int(replicas_to_aggregate)
        self._total_num_replicas = int(total_num_replicas or replicas_to_aggregate)
        self._tokens_per_step = max(self._total_num_replicas, self._replicas_to_aggregate)
        self._gradients_applied = False
        self._global_step: Optional[Variable] = None
        self._local_step: Optional[Variable] = None
        self._vars: List[Variable] = []
        self._acc_names: List[str] = []
        self.chief_init_op: Optional[Op] = None
        self.local_step_init_op: Optional[Op] = None
        self.ready_for_local_init_op: Optional[Tensor] = None
        self._stop_event = threading.Event()
        self._runner: Optional[threading.Thread] = None
        self.steps_applied = 0

    # delegate gradient computation
    def compute_gradients(self, *args, **kwargs):
        return self._opt.compute_gradients(*args, **kwargs)

    def _token_host(self):
        return self._global_step._host()

    def _ensure_accumulators(self):
        for v, an in zip(self._vars, self._acc_names):
            v._host().acc_create(an, list(v.shape), "float32")
        self._token_host().queue_create(self.TOKEN_QUEUE)

    def apply_gradients(self, grads_and_vars, global_step=None, name=None) -> Op:
        if global_step is None:
            raise ValueError("Global step is required to check staleness")
        gv = [(g, v) for g, v in grads_and_vars if g is not None]
        if not gv:
            raise ValueError("Must supply at least one variable")
        self._global_step = global_step
        self._vars = [v for _, v in gv]
        self._acc_names = ["%s/grad_accum" % v.op.name for v in self._vars]
        self._opt._create_slots(self._vars)
        g = fw.get_default_graph()
        saved_scope, g._name_scope = g._name_scope, []
        try:
            with fw.device(None):
                # per-worker LOCAL variable, like [TF] 'sync_rep_local_step'
                self._local_step = Variable(torch.zeros((), dtype=torch.int64), trainable=False,
                                            collections=[GraphKeys.LOCAL_VARIABLES],
                                            name="sync_rep_local_step")
        finally:
            g._name_scope = saved_scope

        def local_init(ctx):
            self._ensure_accumulators()
            self._local_step.assign(global_step.read_value())

        def chief_init(ctx):
            self._ensure_accumulators()
            gs = int(global_step.read_value().item())
            for v, an in zip(self._vars, self._acc_names):
                v._host().acc_set_global_step([an], gs)
            self._local_step.assign(gs)

        self.local_step_init_op = Op(local_init, "sync_replicas/local_step_init")
        self.chief_init_op = Op(chief_init, "sync_replicas/chief_init")
        self.ready_for_local_init_op = fw.report_uninitialized_variables()
        self._gradients_applied = True

        def train(ctx):
            self._ensure_accumulators()
            ls = int(self._local_step.read_value().item())
            by_host: Dict[Tuple, Tuple[list, list, object]] = {}
            for (gnode, v), an in zip(gv, self._acc_names):
                key = (v._device_spec.job, v._device_spec.task)
                names, grads, _ = by_host.setdefault(key, ([], [], v._host()))
                names.append(an)
                grads.append(torch.as_tensor(evaluate(gnode, ctx)).detach().float())
            for names, grads, host in by_host.values():
                host.acc_apply_grad(names, grads, ls)
            token = self._dequeue_token(ctx)
            self._local_step.assign(token)

        return Op(train, name or "sync_replicas/train")

    def _dequeue_token(self, ctx) -> int:
        host = self._token_host()
        sess = getattr(ctx, "session", None)
        while True:
            tok = host.queue_dequeue(self.TOKEN_QUEUE, 0.25)
            if tok is not None:
                return int(tok)
            # queue closed and drained?
            if host.queue_size(self.TOKEN_QUEUE) == 0 and host.kv_get("sync_q_closed", 0.0):
                raise OutOfRangeError("sync_token_q is closed")
            if sess is not None and getattr(sess, "_force_stop", False):
                raise OutOfRangeError("session is stopping")

    # -------------------------------------------------------------- init ops
    def get_init_tokens_op(self, num_tokens: int = -1) -> Op:
        """Tokens that let workers start; ``-1`` -> ``replicas_to_aggregate``
        (Synchronous-SGD-different-learning-rates/ssgd.py:68)."""
        if not self._gradients_applied:
            raise ValueError("get_init_tokens_op() should be called after apply_gradients().")
        tokens_needed = self._replicas_to_aggregate - self._total_num_replicas
        if num_tokens == -1:
            num_tokens = self._replicas_to_aggregate
        elif num_tokens < tokens_needed:
            raise ValueError("Too few tokens to finish the first step: %d (given) vs %d (needed)"
                             % (num_tokens, tokens_needed))

        def run(ctx):
            if num_tokens > 0:
                gs = int(self._global_step.read_value().item())
                self._token_host().queue_enqueue_many(self.TOKEN_QUEUE, [gs] * num_tokens)
        return Op(run, "sync_replicas/init_tokens")

    # ---------------------------------------------------- chief queue runner
    def _aggregate_once(self, timeout: float) -> bool:
        """take_grad(N) -> wrapped optimizer on ps -> global_step += 1 -> tokens."""
        by_host: Dict[Tuple, Tuple[list, list, object]] = {}
        for v, an in zip(self._vars, self._acc_names):
            key = (v._device_spec.job, v._device_spec.task)
            vs, names, _ = by_host.setdefault(key, ([], [], v._host()))
            vs.append(v)
            names.append(an)
        taken = []
        first = True
        for vs, names, host in by_host.values():
            # Only the first take may time out; once it returned, the step is
            # committed and the remaining accumulators are waited on for real.
            grads = host.acc_take_grad(names, self._replicas_to_aggregate,
                                       timeout if first else 30.0)
            if grads is None:
                if first:
                    return False
                raise RuntimeError("accumulators diverged across ps tasks")
            first = False
            taken.append((vs, grads, host))
        for vs, grads, host in taken:
            self._opt._apply_on_host(host, vs, grads)
        gs_host = self._global_step._host()
        new_gs = gs_host.counter_add(self._global_step.storage_key, 1)
        for vs, names, host in by_host.values():
            host.acc_set_global_step(names, new_gs)
        gs_host.queue_enqueue_many(self.TOKEN_QUEUE, [new_gs] * self._tokens_per_step)
        self.steps_applied += 1
        return True

    def _runner_loop(self):
        from ..parallel.rpc import AbortedError
        while not self._stop_event.is_set():
            try:
                self._aggregate_once(timeout=0.2)
            except (AbortedError, KeyError):
                time.sleep(0.05)

    def start_queue_runner(self):
        if self._runner is not None and self._runner.is_alive():
            return
        self._stop_event.clear()
        self._runner = threading.Thread(target=self._runner_loop, daemon=True,
                                        name="sync-replicas-chief")
        self._runner.start()

    def stop_queue_runner(self, close_queue: bool = True):
        self._stop_event.set()
        if self._runner is not None:
            self._runner.join(timeout=5.0)
            self._runner = None
        if close_queue and self._global_step is not None:
            try:
                self._token_host().kv_set("sync_q_closed", "1")
                self._token_host().queue_close(self.TOKEN_QUEUE)
            except Exception:
                pass

    def make_session_run_hook(self, is_chief: bool, num_tokens: int = -1):
        from ..session.hooks import SyncReplicasOptimizerHook
        return SyncReplicasOptimizerHook(self, is_chief, num_tokens)

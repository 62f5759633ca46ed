"""Session run hooks.

The reference's README lists TF's hook vocabulary (README.md:83-92) and the
scripts use ``StopAtStepHook`` (DOWNPOUR/DOWNPOUR.py:116, ADAG/ADAG.py:104,
Synchronous-SGD/ssgd.py:61) and the SyncReplicas hook
(Synchronous-SGD/ssgd.py:60).  ``MonitoredTrainingSession`` adds
``CheckpointSaverHook`` / ``StepCounterHook`` / ``SummarySaverHook`` on the chief
when a ``checkpoint_dir`` is given (DOWNPOUR/DOWNPOUR.py:121-127).

Protocol (same five calls as TF): ``begin()`` once before the session exists,
``after_create_session(session, coord)`` after (re)creation, ``before_run(ctx)``
-> optional ``SessionRunArgs`` with extra fetches, ``after_run(ctx, values)``,
``end(session)`` at clean shutdown.
"""
from __future__ import annotations

import json
import math
import os
import time
from typing import Any, Callable, Dict, List, Optional

import torch

from ..framework import get_global_step

__all__ = [
    "SessionRunHook", "SessionRunArgs", "SessionRunValues", "SessionRunContext",
    "StopAtStepHook", "LoggingTensorHook", "StepCounterHook", "CheckpointSaverHook",
    "NanTensorHook", "NanLossDuringTrainingError", "SummarySaverHook",
    "GlobalStepWaiterHook", "FinalOpsHook", "FeedFnHook", "SyncReplicasOptimizerHook",
    "SecondOrStepTimer", "HeartbeatHook",
]


class SessionRunArgs:
    def __init__(self, fetches=None, feed_dict=None):
        self.fetches, self.feed_dict = fetches, feed_dict


class SessionRunValues:
    def __init__(self, results=None):
        self.results = results


class SessionRunContext:
    def __init__(self, original_args: SessionRunArgs, session):
        self.original_args = original_args
        self.session = session
        self._stop_requested = False

    @property
    def stop_requested(self) -> bool:
        return self._stop_requested

    def request_stop(self) -> None:
        self._stop_requested = True


class SessionRunHook:
    def begin(self): pass
    def after_create_session(self, session, coord): pass
    def before_run(self, run_context: SessionRunContext) -> Optional[SessionRunArgs]: return None
    def after_run(self, run_context: SessionRunContext, run_values: SessionRunValues): pass
    def end(self, session): pass


class SecondOrStepTimer:
    def __init__(self, every_secs: Optional[float] = None, every_steps: Optional[int] = None):
        if (every_secs is None) == (every_steps is None):
            raise ValueError("exactly one of every_secs and every_steps must be set")
        self._secs, self._steps = every_secs, every_steps
        self._last_time: Optional[float] = None
        self._last_step: Optional[int] = None

    def should_trigger_for_step(self, step: int) -> bool:
        if self._last_step is None:
            return True
        if self._last_step == step:
            return False
        if self._secs is not None:
            return time.time() >= self._last_time + self._secs
        return step >= self._last_step + self._steps

    def update_last_triggered_step(self, step: int):
        now = time.time()
        if self._last_time is None:
            dt, ds = None, None
        else:
            dt, ds = now - self._last_time, step - self._last_step
        self._last_time, self._last_step = now, step
        return dt, ds

    def last_triggered_step(self):
        return self._last_step


def _require_global_step():
    gs = get_global_step()
    if gs is None:
        raise RuntimeError("Global step should be created to use this hook.")
    return gs


class StopAtStepHook(SessionRunHook):
    """Request stop once ``global_step >= last_step`` (checked after each run).
    ``num_steps`` counts from the step seen at session creation."""

    def __init__(self, num_steps: Optional[int] = None, last_step: Optional[int] = None):
        if (num_steps is None) == (last_step is None):
            raise ValueError("One of num_steps or last_step must be specified.")
        self._num_steps, self._last_step = num_steps, last_step
        self._gs = None

    def begin(self):
        self._gs = _require_global_step()

    def after_create_session(self, session, coord):
        if self._last_step is None:
            self._last_step = int(self._gs.read_value().item()) + self._num_steps

    def before_run(self, run_context):
        return SessionRunArgs(self._gs)

    def after_run(self, run_context, run_values):
        step = int(run_values.results.item())
        if step >= self._last_step:
            # [TF] re-reads once to skip a stale cached value; reads here are
            # always fresh so one check suffices.
            run_context.request_stop()


class LoggingTensorHook(SessionRunHook):
    def __init__(self, tensors, every_n_iter: Optional[int] = None,
                 every_n_secs: Optional[float] = None, formatter: Optional[Callable] = None,
                 at_end: bool = False, output: Callable[[str], None] = print):
        if isinstance(tensors, dict):
            self._tensors = tensors
        else:
            self._tensors = {getattr(t, "name", str(i)): t for i, t in enumerate(tensors)}
        self._timer = SecondOrStepTimer(every_secs=every_n_secs, every_steps=every_n_iter) \
            if (every_n_iter or every_n_secs) else None
        self._formatter, self._at_end, self._out = formatter, at_end, output
        self._iter = 0
        self.logged: List[Dict[str, Any]] = []

    def begin(self):
        self._iter = 0

    def before_run(self, run_context):
        self._should = self._timer is not None and self._timer.should_trigger_for_step(self._iter)
        return SessionRunArgs(self._tensors) if self._should else None

    def _log(self, vals):
        self.logged.append(vals)
        if self._formatter:
            self._out(self._formatter(vals))
        else:
            self._out(", ".join("%s = %s" % (k, _fmt(v)) for k, v in vals.items()))

    def after_run(self, run_context, run_values):
        if self._should:
            self._log(run_values.results)
            self._timer.update_last_triggered_step(self._iter)
        self._iter += 1

    def end(self, session):
        if self._at_end:
            self._log(session.run(self._tensors))


def _fmt(v):
    if isinstance(v, torch.Tensor):
        return v.tolist()
    return v


class StepCounterHook(SessionRunHook):
    """global_step/sec (and samples/sec when ``batch_size`` is given)."""

    def __init__(self, every_n_steps: int = 100, every_n_secs: Optional[float] = None,
                 output_dir: Optional[str] = None, batch_size: Optional[int] = None,
                 output: Callable[[str], None] = print):
        self._timer = SecondOrStepTimer(every_secs=every_n_secs) if every_n_secs \
            else SecondOrStepTimer(every_steps=every_n_steps)
        self._output_dir, self._batch, self._out = output_dir, batch_size, output
        self.rates: List[float] = []

    def begin(self):
        self._gs = _require_global_step()

    def before_run(self, run_context):
        return SessionRunArgs(self._gs)

    def after_run(self, run_context, run_values):
        step = int(run_values.results.item())
        if self._timer.should_trigger_for_step(step):
            dt, ds = self._timer.update_last_triggered_step(step)
            if dt is not None and dt > 0:
                rate = ds / dt
                self.rates.append(rate)
                msg = "global_step/sec: %g" % rate
                if self._batch:
                    msg += "  samples/sec: %g" % (rate * self._batch)
                self._out(msg)
                if self._output_dir:
                    _append_jsonl(os.path.join(self._output_dir, "metrics.jsonl"),
                                  {"step": step, "global_step/sec": rate, "time": time.time()})


def _append_jsonl(path, obj):
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "a") as f:
        f.write(json.dumps(obj) + "\n")


class CheckpointSaverHook(SessionRunHook):
    """Chief-only periodic save to ``checkpoint_dir/model.ckpt-<step>``."""

    def __init__(self, checkpoint_dir: str, save_secs: Optional[float] = None,
                 save_steps: Optional[int] = None, saver=None,
                 checkpoint_basename: str = "model.ckpt", scaffold=None, listeners=None):
        self._dir = checkpoint_dir
        self._path = os.path.join(checkpoint_dir, checkpoint_basename)
        self._timer = SecondOrStepTimer(every_secs=save_secs, every_steps=save_steps)
        self._saver, self._scaffold = saver, scaffold
        self._listeners = listeners or []
        self.saved_steps: List[int] = []

    def _get_saver(self):
        if self._saver is not None:
            return self._saver
        if self._scaffold is not None and self._scaffold.saver is not None:
            return self._scaffold.saver
        from .checkpoint import Saver
        self._saver = Saver()
        return self._saver

    def begin(self):
        self._gs = _require_global_step()

    def after_create_session(self, session, coord):
        self._save(session, int(self._gs.read_value().item()))

    def before_run(self, run_context):
        return SessionRunArgs(self._gs)

    def after_run(self, run_context, run_values):
        step = int(run_values.results.item())
        if self._timer.should_trigger_for_step(step):
            self._save(run_context.session, step)

    def end(self, session):
        step = int(self._gs.read_value().item())
        if step != self._timer.last_triggered_step():
            self._save(session, step)

    def _save(self, session, step):
        self._timer.update_last_triggered_step(step)
        for l in self._listeners:
            getattr(l, "before_save", lambda *a: None)(session, step)
        self._get_saver().save(session, self._path, global_step=step)
        self.saved_steps.append(step)
        for l in self._listeners:
            getattr(l, "after_save", lambda *a: None)(session, step)


class NanLossDuringTrainingError(RuntimeError):
    def __str__(self):
        return "NaN loss during training."


class NanTensorHook(SessionRunHook):
    def __init__(self, loss_tensor, fail_on_nan_loss: bool = True):
        self._loss, self._fail = loss_tensor, fail_on_nan_loss

    def before_run(self, run_context):
        return SessionRunArgs(self._loss)

    def after_run(self, run_context, run_values):
        v = run_values.results
        v = float(v.float().sum().item()) if isinstance(v, torch.Tensor) else float(v)
        if math.isnan(v):
            if self._fail:
                raise NanLossDuringTrainingError()
            run_context.request_stop()


class SummarySaverHook(SessionRunHook):
    """Scalar summaries to ``<output_dir>/summaries.jsonl`` (the event-file
    role; no TensorBoard dependency)."""

    def __init__(self, save_steps: Optional[int] = None, save_secs: Optional[float] = None,
                 output_dir: Optional[str] = None, summary_op=None, scaffold=None):
        self._timer = SecondOrStepTimer(every_secs=save_secs, every_steps=save_steps)
        self._dir, self._summary = output_dir, summary_op

    def begin(self):
        self._gs = _require_global_step()

    def before_run(self, run_context):
        fetch = {"step": self._gs}
        if self._summary is not None:
            fetch["summary"] = self._summary
        return SessionRunArgs(fetch)

    def after_run(self, run_context, run_values):
        step = int(run_values.results["step"].item())
        if self._summary is None or not self._dir:
            return
        if self._timer.should_trigger_for_step(step):
            self._timer.update_last_triggered_step(step)
            vals = run_values.results["summary"]
            if not isinstance(vals, dict):
                vals = {"value": vals}
            rec = {"step": step, "time": time.time()}
            for k, v in vals.items():
                rec[k] = _fmt(v)
            _append_jsonl(os.path.join(self._dir, "summaries.jsonl"), rec)


class GlobalStepWaiterHook(SessionRunHook):
    """Delay a worker until ``global_step >= wait_until_step``."""

    def __init__(self, wait_until_step: int, poll_secs: float = 0.5):
        self._until, self._poll = int(wait_until_step), poll_secs
        self._started = False

    def begin(self):
        self._gs = _require_global_step()
        self._started = False

    def before_run(self, run_context):
        if self._started or self._until <= 0:
            self._started = True
            return None
        while int(self._gs.read_value().item()) < self._until:
            time.sleep(self._poll)
        self._started = True
        return None


class FinalOpsHook(SessionRunHook):
    def __init__(self, final_ops, final_ops_feed_dict=None):
        self._ops, self._feed = final_ops, final_ops_feed_dict
        self.final_ops_values = None

    def end(self, session):
        if self._ops is not None:
            self.final_ops_values = session.run(self._ops, feed_dict=self._feed)


class FeedFnHook(SessionRunHook):
    def __init__(self, feed_fn: Callable[[], dict]):
        self._feed_fn = feed_fn

    def before_run(self, run_context):
        return SessionRunArgs(fetches=None, feed_dict=self._feed_fn())


class HeartbeatHook(SessionRunHook):
    """Writes this task's liveness word on the ps control plane every run; the
    chief (or a watchdog) reads ``stale_tasks`` to detect dead workers.  Not in
    the reference (it relies on sleeps, DOWNPOUR/DOWNPOUR.py:131,144)."""

    def __init__(self, who: str, host_getter: Callable[[], Any], min_interval: float = 0.5):
        self._who, self._get, self._min = who, host_getter, min_interval
        self._last = 0.0

    def before_run(self, run_context):
        now = time.time()
        if now - self._last >= self._min:
            self._get().heartbeat(self._who)
            self._last = now
        return None


class SyncReplicasOptimizerHook(SessionRunHook):
    """[TF] ``_SyncReplicasOptimizerHook``: after the session exists run
    ``chief_init_op`` (chief) or ``local_step_init_op``; the chief also seeds the
    token queue and starts the aggregation thread (Synchronous-SGD/ssgd.py:60)."""

    def __init__(self, sync_optimizer, is_chief: bool, num_tokens: int = -1):
        self._opt, self._is_chief, self._num_tokens = sync_optimizer, is_chief, num_tokens
        self._init_tokens_op = None

    def begin(self):
        if not self._opt._gradients_applied:
            raise ValueError("SyncReplicasOptimizer.apply_gradient should be called before "
                             "using the hook.")
        if self._is_chief:
            self._local_init_op = self._opt.chief_init_op
            self._init_tokens_op = self._opt.get_init_tokens_op(self._num_tokens)
        else:
            self._local_init_op = self._opt.local_step_init_op

    def after_create_session(self, session, coord):
        session.run(self._local_init_op)
        if self._is_chief:
            session.run(self._init_tokens_op)
            self._opt.start_queue_runner()

    def end(self, session):
        if self._is_chief:
            self._opt.stop_queue_runner(close_queue=True)

"""Session lifecycle: ``Session``, ``Scaffold``, ``MonitoredTrainingSession`` and
``Supervisor``.

Reference call sites: ``tf.Session(target)`` + ``list_devices``
(Basics-Tutorial/Servers.ipynb:180,251); ``MonitoredTrainingSession(master,
is_chief, config, scaffold, hooks, save_checkpoint_secs, checkpoint_dir,
stop_grace_period_secs)`` (DOWNPOUR/DOWNPOUR.py:121-127,
Synchronous-SGD/ssgd.py:65-69); ``Scaffold(init_op, local_init_op,
ready_for_local_init_op)`` (DOWNPOUR/DOWNPOUR.py:118,
Synchronous-SGD-different-learning-rates/ssgd.py:81-83); ``Supervisor(logdir,
is_chief, save_model_secs)`` + ``prepare_or_wait_for_session`` +
``should_stop`` (Hogwild/Hogwild.py:47-52,
Distributed-Setup/dist_setup_sup.py:43-46).

Bring-up protocol reproduced ([TF] behaviour):

* chief  : restore the latest checkpoint of ``checkpoint_dir`` if there is one,
  else run ``scaffold.init_op``; then ``local_init_op``.
* others : poll ``ready_op`` (no uninitialised GLOBAL variable) and only then
  run ``local_init_op``.
* every ``run``: ``before_run`` of each hook -> evaluate -> ``after_run``;
  ``AbortedError`` (a peer went away) re-creates the session through the same
  path, so a restarted chief resumes from its last checkpoint.
* ``close()``: hooks' ``end`` then wait at most ``stop_grace_period_secs``.
"""
from __future__ import annotations

import os
import threading
import time
from typing import Callable, List, Optional, Sequence

import torch

from .. import framework as fw
from .. import runtime
from ..framework import RunContext, Variable, evaluate
from ..parallel.rpc import AbortedError
from ..framework import OutOfRangeError, get_global_step
from . import hooks as hooks_lib
from .checkpoint import Saver, latest_checkpoint
from .hooks import (SessionRunArgs, SessionRunContext, SessionRunHook, SessionRunValues)

__all__ = ["Session", "Scaffold", "MonitoredSession", "MonitoredTrainingSession",
           "Supervisor", "Coordinator", "SessionManager", "SessionCreator",
           "ChiefSessionCreator", "WorkerSessionCreator"]


class Coordinator:
    """Minimal [TF] Coordinator: a stop flag shared by background threads."""

    def __init__(self):
        self._ev = threading.Event()
        self._threads: List[threading.Thread] = []

    def request_stop(self, ex=None):
        self._ev.set()

    def should_stop(self) -> bool:
        return self._ev.is_set()

    def clear_stop(self):
        self._ev.clear()

    def register_thread(self, t):
        self._threads.append(t)

    def join(self, threads=None, stop_grace_period_secs: float = 120):
        deadline = time.time() + stop_grace_period_secs
        for t in list(threads or self._threads):
            t.join(timeout=max(0.0, deadline - time.time()))

    def wait_for_stop(self, timeout=None) -> bool:
        return self._ev.wait(timeout)


class Session:
    """Evaluates fetches against the cluster this process's ``Server`` belongs to.

    Fetches may be ``Tensor`` / ``Op`` nodes, ``Variable`` s (their current
    value), plain callables, or nested lists/tuples/dicts of those.  They are
    evaluated **in the order given** within one memoised ``RunContext`` -- a
    deterministic version of TF's unordered fetch.
    """

    def __init__(self, target: str = "", config=None):
        self.target = target
        self.config = config
        self._closed = False
        self._force_stop = False
        self._server = runtime.current_server()
        # ConfigProto(log_device_placement=True): the first time an op runs in this session,
        # print "<name>: (<type>): <device>" (TF's device-mapping log; the reference switches
        # it off explicitly, DOWNPOUR/DOWNPOUR.py:21)
        self._log_placement = bool(getattr(config, "log_device_placement", False))
        self._placement_logged = set()

    def _log_node(self, node) -> None:
        if node.op.name not in self._placement_logged:
            self._placement_logged.add(node.op.name)
            print("%s: (%s): %s" % (node.op.name, node.type, node.device or "<unplaced>"))

    def run(self, fetches, feed_dict: Optional[dict] = None):
        if self._closed:
            raise RuntimeError("Attempted to use a closed Session.")
        ctx = RunContext(self)
        if feed_dict:
            for k, v in feed_dict.items():
                ctx.feed[id(k)] = torch.as_tensor(v)
                if isinstance(k, fw.Node):
                    ctx.cache[id(k)] = torch.as_tensor(v)
        return self._fetch(fetches, ctx)

    def _fetch(self, f, ctx):
        if f is None:
            return None
        if isinstance(f, (fw.Node, Variable)):
            return evaluate(f, ctx)
        if isinstance(f, (list, tuple)):
            return type(f)(self._fetch(x, ctx) for x in f) if not hasattr(f, "_fields") \
                else type(f)(*[self._fetch(x, ctx) for x in f])
        if isinstance(f, dict):
            return {k: self._fetch(v, ctx) for k, v in f.items()}
        if callable(f):
            return f()
        return f

    def list_devices(self):
        return self._server.list_devices()

    def close(self):
        self._closed = True

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


class Scaffold:
    """Holds the three init stages (DOWNPOUR/DOWNPOUR.py:118)."""

    def __init__(self, init_op=None, init_feed_dict=None, init_fn: Optional[Callable] = None,
                 ready_op=None, ready_for_local_init_op=None, local_init_op=None,
                 summary_op=None, saver: Optional[Saver] = None):
        self.init_op, self.init_feed_dict, self.init_fn = init_op, init_feed_dict, init_fn
        self.ready_op, self.ready_for_local_init_op = ready_op, ready_for_local_init_op
        self.local_init_op, self.summary_op, self.saver = local_init_op, summary_op, saver
        self._finalized = False

    def finalize(self) -> "Scaffold":
        if self._finalized:
            return self
        if self.init_op is None:
            self.init_op = fw.global_variables_initializer()
        if self.ready_op is None:
            self.ready_op = fw.report_uninitialized_variables()
        if self.ready_for_local_init_op is None:
            self.ready_for_local_init_op = fw.report_uninitialized_variables()
        if self.local_init_op is None:
            self.local_init_op = fw.local_variables_initializer()
        if self.saver is None:
            self.saver = Saver()
        self._finalized = True
        return self


class SessionManager:
    """[TF] SessionManager: chief ``prepare_session`` / worker ``wait_for_session``."""

    def __init__(self, scaffold: Scaffold, recovery_wait_secs: float = 0.2):
        self._sc = scaffold.finalize()
        self._wait = recovery_wait_secs

    @staticmethod
    def _run_local_init(sess, op):
        if isinstance(op, (list, tuple)):
            for o in op:
                sess.run(o)
        elif op is not None:
            sess.run(op)

    def prepare_session(self, master, checkpoint_dir=None, config=None) -> Session:
        sess = Session(master, config)
        restored = False
        if checkpoint_dir:
            ck = latest_checkpoint(checkpoint_dir)
            if ck:
                # Variables absent from the checkpoint keep their init value.
                sess.run(self._sc.init_op, feed_dict=self._sc.init_feed_dict)
                self._sc.saver.restore(sess, ck)
                restored = True
        if not restored:
            sess.run(self._sc.init_op, feed_dict=self._sc.init_feed_dict)
            if self._sc.init_fn is not None:
                self._sc.init_fn(self._sc, sess)
        self._run_local_init(sess, self._sc.local_init_op)
        not_ready = sess.run(self._sc.ready_op)
        if not_ready:
            raise RuntimeError("Init operations did not make model ready. Uninitialized: %s"
                               % (not_ready,))
        self.restored = restored
        return sess

    def wait_for_session(self, master, config=None, max_wait_secs: float = float("inf")) -> Session:
        sess = Session(master, config)
        t0 = time.time()
        while True:
            try:
                if not sess.run(self._sc.ready_for_local_init_op):
                    self._run_local_init(sess, self._sc.local_init_op)
                    if not sess.run(self._sc.ready_op):
                        return sess
            except AbortedError:
                pass
            if time.time() - t0 > max_wait_secs:
                raise TimeoutError("Session was not ready after waiting %d secs." % max_wait_secs)
            time.sleep(self._wait)


class SessionCreator:
    """[TF] ``tf.train.SessionCreator`` ("Factory for creating a session?", README.md:77): the
    object a ``MonitoredSession`` asks for a session, at start-up and again after every
    ``AbortedError`` recovery."""

    def create_session(self) -> Session:
        raise NotImplementedError("create_session is not implemented for %s" % type(self).__name__)


class ChiefSessionCreator(SessionCreator):
    """Chief path: restore the latest checkpoint of ``checkpoint_dir`` (or the explicit
    ``checkpoint_filename_with_path``), else run the scaffold's init stages."""

    def __init__(self, scaffold: Optional[Scaffold] = None, master: str = "", config=None,
                 checkpoint_dir: Optional[str] = None,
                 checkpoint_filename_with_path: Optional[str] = None,
                 recovery_wait_secs: float = 0.2):
        self._scaffold = scaffold or Scaffold()
        self._master, self._config = master, config
        self._checkpoint_dir = checkpoint_dir
        self._checkpoint_file = checkpoint_filename_with_path
        self._recovery_wait = recovery_wait_secs
        self.restored = False

    @property
    def scaffold(self) -> Scaffold:
        return self._scaffold

    def create_session(self) -> Session:
        sm = SessionManager(self._scaffold.finalize(), self._recovery_wait)
        if self._checkpoint_file:
            sess = sm.prepare_session(self._master, None, self._config)
            self._scaffold.saver.restore(sess, self._checkpoint_file)
            self.restored = True
            return sess
        sess = sm.prepare_session(self._master, self._checkpoint_dir, self._config)
        self.restored = sm.restored
        return sess


class WorkerSessionCreator(SessionCreator):
    """Non-chief path: poll until the chief has initialised the GLOBAL variables, then run the
    local init stage (TF default ``max_wait_secs`` 30 min)."""

    def __init__(self, scaffold: Optional[Scaffold] = None, master: str = "", config=None,
                 max_wait_secs: float = 30 * 60, recovery_wait_secs: float = 0.2):
        self._scaffold = scaffold or Scaffold()
        self._master, self._config = master, config
        self._max_wait, self._recovery_wait = max_wait_secs, recovery_wait_secs

    @property
    def scaffold(self) -> Scaffold:
        return self._scaffold

    def create_session(self) -> Session:
        sm = SessionManager(self._scaffold.finalize(), self._recovery_wait)
        return sm.wait_for_session(self._master, self._config, self._max_wait)


class MonitoredSession:
    """Hooked session with recovery.  ``should_stop()`` / ``run()`` / ``close()``.

    Either TF's form ``MonitoredSession(session_creator=ChiefSessionCreator(...), hooks=...)``
    or the keyword form ``MonitoredTrainingSession`` uses (``master`` / ``is_chief`` /
    ``checkpoint_dir``), from which the matching creator is built."""

    def __init__(self, master: str = "", is_chief: bool = True, scaffold: Optional[Scaffold] = None,
                 hooks: Optional[Sequence[SessionRunHook]] = None, checkpoint_dir: Optional[str] = None,
                 config=None, stop_grace_period_secs: float = 120, max_wait_secs: float = 7200,
                 recovery_wait_secs: float = 0.2,
                 session_creator: Optional[SessionCreator] = None):
        if isinstance(master, SessionCreator):           # TF's first positional argument
            session_creator, master = master, ""
        if session_creator is not None:
            scaffold = getattr(session_creator, "scaffold", None) or scaffold
            is_chief = not isinstance(session_creator, WorkerSessionCreator)
        self._master, self._is_chief = master, is_chief
        self._scaffold = (scaffold or Scaffold())
        self._hooks = list(hooks or [])
        self._checkpoint_dir, self._config = checkpoint_dir, config
        self._grace = stop_grace_period_secs
        self._max_wait, self._recovery_wait = max_wait_secs, recovery_wait_secs
        if session_creator is None:
            session_creator = (
                ChiefSessionCreator(self._scaffold, master, config, checkpoint_dir,
                                    recovery_wait_secs=recovery_wait_secs) if is_chief else
                WorkerSessionCreator(self._scaffold, master, config, max_wait_secs,
                                     recovery_wait_secs))
        self._creator = session_creator
        self._coord = Coordinator()
        self._should_stop = False
        self._closed = False
        self._force_stop = False
        self.num_recoveries = 0
        for h in self._hooks:
            h.begin()
        self._scaffold.finalize()
        self._sess = self._create_session()

    # ------------------------------------------------------------------ setup
    def _create_session(self) -> Session:
        while True:
            try:
                sess = self._creator.create_session()
                for h in self._hooks:
                    h.after_create_session(sess, self._coord)
                return sess
            except AbortedError:
                time.sleep(self._recovery_wait)

    # -------------------------------------------------------------------- run
    def run(self, fetches, feed_dict: Optional[dict] = None):
        while True:
            try:
                return self._run_once(fetches, feed_dict)
            except AbortedError:
                # "initialization, hooks, recovery" (README.md:78)
                self.num_recoveries += 1
                self._sess.close()
                self._sess = self._create_session()
            except OutOfRangeError:
                self._should_stop = True
                return None

    def _run_once(self, fetches, feed_dict):
        rc = SessionRunContext(SessionRunArgs(fetches, feed_dict), self._sess)
        extra, feeds = [], dict(feed_dict or {})
        for h in self._hooks:
            req = h.before_run(rc)
            extra.append(req.fetches if req is not None else None)
            if req is not None and req.feed_dict:
                feeds.update(req.feed_dict)
        self._sess._force_stop = self._force_stop
        user, hook_vals = self._sess.run([fetches, extra], feed_dict=feeds or None)
        for h, v in zip(self._hooks, hook_vals):
            h.after_run(rc, SessionRunValues(v))
        if rc.stop_requested:
            self._should_stop = True
        return user

    def should_stop(self) -> bool:
        return self._should_stop or self._closed or self._coord.should_stop()

    def request_stop(self):
        self._should_stop = True

    @property
    def raw_session(self) -> Session:
        return self._sess

    def list_devices(self):
        return self._sess.list_devices()

    # ------------------------------------------------------------------ close
    def close(self):
        if self._closed:
            return
        self._force_stop = True
        try:
            for h in self._hooks:
                try:
                    h.end(self._sess)
                except AbortedError:
                    pass
        finally:
            self._coord.request_stop()
            self._coord.join(stop_grace_period_secs=self._grace)
            self._sess.close()
            self._closed = True
            self._should_stop = True

    def __enter__(self):
        return self

    def __exit__(self, et, ev, tb):
        self.close()
        return et is not None and issubclass(et, OutOfRangeError)


def MonitoredTrainingSession(master: str = "", is_chief: bool = True,
                             checkpoint_dir: Optional[str] = None,
                             scaffold: Optional[Scaffold] = None,
                             hooks: Optional[Sequence[SessionRunHook]] = None,
                             chief_only_hooks: Optional[Sequence[SessionRunHook]] = None,
                             save_checkpoint_secs: Optional[float] = 600,
                             save_checkpoint_steps: Optional[int] = None,
                             save_summaries_steps: Optional[int] = 100,
                             save_summaries_secs: Optional[float] = None,
                             config=None, stop_grace_period_secs: float = 120,
                             log_step_count_steps: Optional[int] = 100,
                             max_wait_secs: float = 7200,
                             recovery_wait_secs: float = 0.2) -> MonitoredSession:
    """Factory with TF's defaults.  With a ``checkpoint_dir`` the chief gets a
    ``CheckpointSaverHook``, a ``StepCounterHook`` and a ``SummarySaverHook``
    (DOWNPOUR/DOWNPOUR.py:121-127: ``save_checkpoint_secs=1,
    checkpoint_dir='logdir'``)."""
    scaffold = scaffold or Scaffold()
    all_hooks: List[SessionRunHook] = []
    if is_chief and chief_only_hooks:
        all_hooks.extend(chief_only_hooks)
    if is_chief and checkpoint_dir and get_global_step() is not None:
        if log_step_count_steps:
            all_hooks.append(hooks_lib.StepCounterHook(
                every_n_steps=log_step_count_steps, output_dir=checkpoint_dir,
                output=lambda s: None))
        if scaffold.summary_op is not None and (save_summaries_steps or save_summaries_secs):
            all_hooks.append(hooks_lib.SummarySaverHook(
                save_steps=None if save_summaries_secs else save_summaries_steps,
                save_secs=save_summaries_secs, output_dir=checkpoint_dir,
                summary_op=scaffold.summary_op))
        if save_checkpoint_steps:
            all_hooks.append(hooks_lib.CheckpointSaverHook(
                checkpoint_dir, save_steps=save_checkpoint_steps, scaffold=scaffold))
        elif save_checkpoint_secs:
            all_hooks.append(hooks_lib.CheckpointSaverHook(
                checkpoint_dir, save_secs=save_checkpoint_secs, scaffold=scaffold))
    if hooks:
        all_hooks.extend(hooks)
    return MonitoredSession(master=master, is_chief=is_chief, scaffold=scaffold, hooks=all_hooks,
                            checkpoint_dir=checkpoint_dir, config=config,
                            stop_grace_period_secs=stop_grace_period_secs,
                            max_wait_secs=max_wait_secs, recovery_wait_secs=recovery_wait_secs)


class Supervisor:
    """Older lifecycle helper used by the Hogwild / Distributed-Setup scripts.

    ``Supervisor()`` with no arguments (Non-Distributed_Setup.py:23) just
    initialises.  With ``logdir`` the chief restores from / saves to
    ``logdir/model.ckpt`` every ``save_model_secs`` on a background thread
    (Hogwild/Hogwild.py:47-49); non-chiefs poll until the model is ready
    (every 30 s in TF; ``recovery_wait_secs`` here defaults to 1 s).
    """

    USE_DEFAULT = object()

    def __init__(self, is_chief: bool = True, logdir: Optional[str] = None, init_op=USE_DEFAULT,
                 local_init_op=USE_DEFAULT, ready_op=USE_DEFAULT,
                 ready_for_local_init_op=USE_DEFAULT, saver=USE_DEFAULT, global_step=USE_DEFAULT,
                 save_model_secs: float = 600, recovery_wait_secs: float = 1.0,
                 stop_grace_secs: float = 120, checkpoint_basename: str = "model.ckpt",
                 init_fn=None, init_feed_dict=None, save_summaries_secs: float = 120):
        D = Supervisor.USE_DEFAULT
        self._is_chief, self._logdir = is_chief, logdir
        self._scaffold = Scaffold(
            init_op=None if init_op is D else init_op,
            init_feed_dict=init_feed_dict, init_fn=(lambda sc, s: init_fn(s)) if init_fn else None,
            ready_op=None if ready_op is D else ready_op,
            ready_for_local_init_op=None if ready_for_local_init_op is D else ready_for_local_init_op,
            local_init_op=None if local_init_op is D else local_init_op,
            saver=None if saver is D else saver).finalize()
        self._global_step = get_global_step() if global_step is D else global_step
        self._save_model_secs = save_model_secs
        self._save_summaries_secs = save_summaries_secs
        self.step_rates: List[float] = []
        self._recovery_wait_secs = recovery_wait_secs
        self._stop_grace_secs = stop_grace_secs
        self._save_path = os.path.join(logdir, checkpoint_basename) if logdir else None
        self._coord = Coordinator()
        self._sess: Optional[Session] = None
        self._threads: List[threading.Thread] = []
        self.saves = 0

    @property
    def saver(self) -> Saver:
        return self._scaffold.saver

    @property
    def coord(self) -> Coordinator:
        return self._coord

    @property
    def is_chief(self) -> bool:
        return self._is_chief

    @property
    def save_path(self):
        return self._save_path

    def prepare_or_wait_for_session(self, master: str = "", config=None,
                                    wait_for_checkpoint=False, max_wait_secs: float = 7200,
                                    start_standard_services: bool = True) -> Session:
        self._coord.clear_stop()
        sm = SessionManager(self._scaffold, self._recovery_wait_secs)
        if self._is_chief:
            self._sess = sm.prepare_session(master, self._logdir, config)
            if start_standard_services:
                self.start_standard_services(self._sess)
        else:
            self._sess = sm.wait_for_session(master, config, max_wait_secs)
        return self._sess

    def start_standard_services(self, sess: Session):
        """[TF] the chief's background services: the timed checkpoint thread
        (``save_model_secs``) and the step counter that logs ``global_step/sec`` every
        ``save_summaries_secs`` to ``logdir/metrics.jsonl`` (TF's ``SVStepCounterThread``)."""
        started = []
        if not (self._is_chief and self._logdir):
            return started
        if self._save_model_secs:
            started.append(threading.Thread(target=self._saver_loop, args=(sess,), daemon=True,
                                            name="supervisor-saver"))
        if self._save_summaries_secs and self._global_step is not None:
            started.append(threading.Thread(target=self._step_counter_loop, args=(sess,),
                                            daemon=True, name="supervisor-step-counter"))
        for t in started:
            t.start()
            self._threads.append(t)
            self._coord.register_thread(t)
        return started

    def _step_counter_loop(self, sess):
        last_t, last_step = time.time(), None
        while not self._coord.wait_for_stop(self._save_summaries_secs):
            try:
                step = int(sess.run(self._global_step))
            except Exception:
                continue
            now = time.time()
            if last_step is not None and now > last_t:
                rate = (step - last_step) / (now - last_t)
                self.step_rates.append(rate)
                hooks_lib._append_jsonl(os.path.join(self._logdir, "metrics.jsonl"),
                                        {"step": step, "global_step/sec": rate, "time": now})
            last_t, last_step = now, step

    def _save(self, sess):
        gs = self._global_step
        self.saver.save(sess, self._save_path, global_step=gs)
        self.saves += 1

    def _saver_loop(self, sess):
        while not self._coord.wait_for_stop(self._save_model_secs):
            try:
                self._save(sess)
            except Exception:
                pass

    def should_stop(self) -> bool:
        return self._coord.should_stop()

    def request_stop(self, ex=None):
        self._coord.request_stop(ex)

    def stop(self, threads=None, close_summary_writer=True):
        self._coord.request_stop()
        self._coord.join(threads, stop_grace_period_secs=self._stop_grace_secs)
        if self._is_chief and self._logdir and self._sess is not None and not self._sess._closed:
            try:
                self._save(self._sess)   # final checkpoint, like managed_session's exit
            except Exception:
                pass

    class _Managed:
        def __init__(self, sv, master, config):
            self._sv, self._master, self._config = sv, master, config

        def __enter__(self):
            return self._sv.prepare_or_wait_for_session(self._master, self._config)

        def __exit__(self, et, ev, tb):
            self._sv.stop()
            if self._sv._sess is not None:
                self._sv._sess.close()
            return False

    def managed_session(self, master: str = "", config=None):
        return Supervisor._Managed(self, master, config)

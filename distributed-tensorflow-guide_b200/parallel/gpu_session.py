"""Session lifecycle for the B200 engines: the same hooks, stop conditions and
checkpoints-by-name as on the host backend, driven by device-resident state.

On CPU the ``global_step`` variable and the model variables live in a ``VariableHost`` and
hooks read them through ``Variable.read_value``.  On B200 they live in the ps shards
(``PeerStore``) or in the sync engine's master copy; ``DeviceBackedVariable`` gives hooks
and the ``Saver`` the same ``Variable`` surface over that memory (a D2H read of 8 bytes for
the step; a gather of the shards for a checkpoint, K15).

Reference flow reproduced (DOWNPOUR/DOWNPOUR.py:116-146): chief restores the latest
checkpoint or seeds the ps, everybody pulls, loop ``while not sess.should_stop()``,
``StopAtStepHook`` on the shared ``global_step``, chief-only timed checkpoints, clean close.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import torch

from .. import framework as fw
from ..framework import GraphKeys, Variable
from ..session import hooks as hooks_lib
from ..session.checkpoint import Saver, latest_checkpoint, load_checkpoint
from ..session.monitored import MonitoredTrainingSession, Scaffold
from .gpu_engines import GpuWorker

__all__ = ["DeviceBackedVariable", "GpuTrainingSession"]


class DeviceBackedVariable(Variable):
    """``Variable`` whose value is produced by a callable (device state) instead of a host."""

    def __init__(self, name: str, reader: Callable[[], torch.Tensor], dtype=torch.int64, shape=()):
        g = fw.get_default_graph()
        self._graph = g
        self._reader = reader
        self._init_fn = None
        self._initial_value = torch.zeros(shape, dtype=dtype)
        self.dtype, self.shape, self.trainable = dtype, tuple(shape), False
        self.op = fw._OpName(g.unique_name(name))
        self.collections = [GraphKeys.GLOBAL_VARIABLES]
        g.add_to_collection(GraphKeys.GLOBAL_VARIABLES, self)
        g._variables_by_name[self.op.name] = self
        from ..cluster import DeviceSpec
        self._device_spec = DeviceSpec(job="ps", task=0, device_type="GPU", device_index=0)
        self.initializer = fw.no_op(self.op.name + "/Assign")

    def _read(self) -> torch.Tensor:
        return torch.as_tensor(self._reader(), dtype=self.dtype)

    def _assign_raw(self, value) -> None:       # device state is owned by the engine
        pass

    def is_initialized(self) -> bool:
        return True


class GpuTrainingSession:
    """``with GpuTrainingSession(worker, ...) as sess: while not sess.should_stop(): sess.step(x, y)``"""

    def __init__(self, worker: GpuWorker, last_step: Optional[int] = None,
                 checkpoint_dir: Optional[str] = None, save_checkpoint_secs: Optional[float] = None,
                 save_checkpoint_steps: Optional[int] = None,
                 hooks: Sequence[hooks_lib.SessionRunHook] = (), step_check_every: int = 1,
                 watchdog_secs: Optional[float] = None, stop_on_stale: bool = False):
        self.worker = worker
        self._check_every = max(1, step_check_every)
        fw.reset_default_graph()
        # Reading the ps-resident step is a D2H copy + stream synchronise; with
        # step_check_every = k the hooks see a value refreshed every k-th run (StopAtStep and
        # the checkpoint cadence may then act up to k-1 steps late, never early).
        self._runs = 0
        self._cached_step: Optional[int] = None

        def read_step():
            if self._cached_step is None or self._runs % self._check_every == 0:
                self._cached_step = int(worker.global_step())
            return self._cached_step
        self.global_step = DeviceBackedVariable("global_step", read_step)
        self._loss = None
        self.loss = fw.tensor(lambda: self._loss, "loss")
        restored = self._maybe_restore(checkpoint_dir)
        self.restored_from = restored
        all_hooks: List[hooks_lib.SessionRunHook] = list(hooks)
        self.heartbeat = self.watchdog = None
        if watchdog_secs is not None:
            # a heartbeat word per rank in peer-mapped HBM, bumped on the training stream after
            # every step; the chief's watchdog names the ranks that went quiet (heartbeat.py)
            from .heartbeat import DeviceHeartbeat, DeviceHeartbeatHook, Watchdog
            self.heartbeat = DeviceHeartbeat(worker.fabric)
            if worker.is_chief:
                self.watchdog = Watchdog(self.heartbeat.read, timeout_secs=watchdog_secs)
            all_hooks.append(DeviceHeartbeatHook(self.heartbeat, self.watchdog,
                                                 stop_on_stale=stop_on_stale))
        if last_step is not None:
            all_hooks.append(hooks_lib.StopAtStepHook(last_step=last_step))
        saver = Saver(var_list=[], extra_state_fn=worker.named_state)
        self._sess = MonitoredTrainingSession(
            is_chief=worker.is_chief, checkpoint_dir=checkpoint_dir if worker.is_chief else None,
            scaffold=Scaffold(init_op=fw.no_op(), local_init_op=fw.no_op(), ready_op=fw.tensor(lambda: []),
                              ready_for_local_init_op=fw.tensor(lambda: []), saver=saver),
            hooks=all_hooks, save_checkpoint_secs=save_checkpoint_secs,
            save_checkpoint_steps=save_checkpoint_steps, log_step_count_steps=None)
        self._train = fw.Op(lambda ctx: self._do_step(), "gpu_train_step")
        self._batch = None

    def _maybe_restore(self, checkpoint_dir) -> Optional[str]:
        """Chief-side resume: ps shards / master copy <- latest checkpoint."""
        w = self.worker
        ck = latest_checkpoint(checkpoint_dir) if checkpoint_dir else None
        if ck is None:
            return None
        data = load_checkpoint(ck)
        if w.store is not None:
            w.store.restore_named(data)
            torch.cuda.synchronize()
            w.fabric.host_barrier()
            w.store.pull(w.model.p32, w.model.p16)
        else:
            full = w.model.p32.clone()
            w.model.layout.pack(full, {k: v for k, v in data.items() if k in w.model.layout.entries})
            w.sync.load_master(full)
            w.sync.step_no = int(data.get("global_step", 0))
            torch.cuda.synchronize()
            w.fabric.host_barrier()
        return ck

    def _do_step(self):
        x, y = self._batch
        self._loss = self.worker.step(x, y)
        self._runs += 1

    def step(self, x, y):
        self._batch = (x, y)
        self._sess.run(self._train)
        return self._loss

    def should_stop(self) -> bool:
        return self._sess.should_stop()

    def close(self):
        self._sess.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False

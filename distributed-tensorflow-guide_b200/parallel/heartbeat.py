"""Failure detection on the device: a heartbeat word per rank in peer-mapped HBM and a
host-side watchdog ("heartbeat word per role in peer memory + host watchdog").

The reference has no failure detection of its own; what TF gives it is session recovery on
``Aborted`` / ``Unavailable`` (/root/reference/README.md:78), SyncReplicas' tolerance of
stragglers when ``replicas_to_aggregate < total_num_replicas``
(/root/reference/Synchronous-SGD/README.md:3) and a ps that has to be killed by hand
(/root/reference/README.md:55-59).  On the host backend ``HeartbeatHook`` +
``VariableHost.stale_tasks`` cover this (``parallel/rpc.py``).  On B200 the data plane has no host
in it, so liveness is read where the work happens:

* ``DeviceHeartbeat``: every rank owns one 64-byte slot ``[beats, done]`` in a symmetric buffer
  every other rank has mapped.  ``beat()`` is a one-thread kernel on the *training stream*
  (``dtg_counter_add``): the word moves only when the GPU has retired the step before it, so a
  rank whose kernels hang looks dead even though its process is alive.  ``read()`` copies every
  rank's slot to pinned host memory on a stream of its own -- it never queues behind training
  kernels, and it keeps working for a rank whose process has exited, because an imported IPC
  mapping outlives its exporter.
* ``Watchdog``: pure host logic over any ``read()``-like callable; a rank is *stale* when its
  counter has not moved for ``timeout_secs`` and it has not marked itself done.  ``poll()`` from
  a hook, or ``start()`` for a background thread.
* ``DeviceHeartbeatHook``: beats after every run, lets the chief poll, marks done at ``end()``.

Asynchronous schemes keep running when a worker dies (nothing waits for it); the synchronous
kernel's N-of-M gate proceeds without a late rank and reports a time-out status for an absent
one (``ops/csrc/sync_kernels.cu``) -- the watchdog says *which* rank that was.
"""
from __future__ import annotations

import threading
import time
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch

from ..session.hooks import SessionRunHook

__all__ = ["DeviceHeartbeat", "Watchdog", "DeviceHeartbeatHook"]


class DeviceHeartbeat:
    """Collective constructor (it allocates a symmetric buffer): every rank creates it at the
    same point of the program."""

    SLOT_BYTES = 64

    def __init__(self, fabric, name: str = "heartbeat"):
        from .. import ops
        self._ops = ops
        self.fabric = fabric
        self.buf = fabric.alloc(name, 256)
        self._slots = [self.buf.tensor(r, torch.int64, 0, 2) for r in range(fabric.world)]
        self._mine = self._slots[fabric.rank]
        self._mine.zero_()
        self._host = torch.zeros(fabric.world, 2, dtype=torch.int64).pin_memory()
        self._stream = torch.cuda.Stream(device=fabric.device)
        torch.cuda.synchronize(fabric.device)
        fabric.host_barrier()                     # every slot zeroed before anyone reads

    def beat(self, n: int = 1) -> None:
        """Stream-ordered: enqueued behind the step it vouches for."""
        self._ops.counter_add(self._mine[:1], n)

    def mark_done(self) -> None:
        self._mine[1:2].fill_(1)

    def read(self) -> List[Tuple[int, int]]:
        """``[(beats, done)] * world`` as of now (peer reads on a private stream)."""
        lib = self._ops.lib()
        with torch.cuda.device(self.fabric.device):
            for r, t in enumerate(self._slots):
                lib.dtg_memcpy_async(self._host[r].data_ptr(), t.data_ptr(), 16,
                                     self._stream.cuda_stream)
            self._stream.synchronize()
        return [(int(b), int(d)) for b, d in self._host.tolist()]


class Watchdog:
    """Host watchdog over heartbeat counters.  ``reader()`` returns ``[(beats, done), ...]``."""

    def __init__(self, reader: Callable[[], Sequence[Tuple[int, int]]], timeout_secs: float = 30.0,
                 on_stale: Optional[Callable[[List[int]], None]] = None,
                 clock: Callable[[], float] = time.time):
        self._reader, self.timeout_secs, self._on_stale, self._clock = reader, timeout_secs, on_stale, clock
        self._last: Dict[int, Tuple[int, float]] = {}
        self.stale: List[int] = []
        self.done: List[int] = []
        self.polls = 0
        self._thread: Optional[threading.Thread] = None
        self._stop = threading.Event()

    def poll(self) -> List[int]:
        now = self._clock()
        stale, done = [], []
        for r, (beats, fin) in enumerate(self._reader()):
            prev = self._last.get(r)
            if prev is None or beats != prev[0]:
                self._last[r] = (beats, now)
            if fin:
                done.append(r)
            elif now - self._last[r][1] > self.timeout_secs:
                stale.append(r)
        newly = [r for r in stale if r not in self.stale]
        self.stale, self.done = stale, done
        self.polls += 1
        if newly and self._on_stale is not None:
            self._on_stale(newly)
        return stale

    def alive(self) -> List[int]:
        return [r for r in self._last if r not in self.stale and r not in self.done]

    # optional background thread (the chief's "host watchdog")
    def start(self, interval_secs: float = 1.0, device=None) -> "Watchdog":
        def loop():
            if device is not None and torch.cuda.is_available():
                torch.cuda.set_device(device)
            while not self._stop.wait(interval_secs):
                try:
                    self.poll()
                except Exception:          # a watchdog never takes training down
                    pass
        self._stop.clear()
        self._thread = threading.Thread(target=loop, daemon=True, name="dtg-watchdog")
        self._thread.start()
        return self

    def stop(self) -> None:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=5.0)
            self._thread = None


class DeviceHeartbeatHook(SessionRunHook):
    """Beat after every run; on the chief poll the watchdog every ``poll_every`` runs and, with
    ``stop_on_stale``, request a stop when a rank has gone quiet."""

    def __init__(self, heartbeat, watchdog: Optional[Watchdog] = None, poll_every: int = 10,
                 stop_on_stale: bool = False, log: Optional[Callable[[str], None]] = print):
        self._hb, self._wd, self._every = heartbeat, watchdog, max(1, poll_every)
        self._stop_on_stale, self._log = stop_on_stale, log
        self._runs = 0
        self.reported: List[int] = []

    def after_run(self, run_context, run_values):
        self._hb.beat()
        self._runs += 1
        if self._wd is None or self._runs % self._every:
            return
        stale = self._wd.poll()
        new = [r for r in stale if r not in self.reported]
        if new:
            self.reported.extend(new)
            if self._log:
                self._log("watchdog: no heartbeat from rank(s) %s for %.1f s" % (
                    new, self._wd.timeout_secs))
        if stale and self._stop_on_stale:
            run_context.request_stop()

    def end(self, session):
        self._hb.mark_done()

"""Variable host + its TCP service/client: the CPU transport and, on B200, the
control plane.

What the reference delegates to TF's gRPC runtime (``protocol: "grpc"``,
Basics-Tutorial/Servers.ipynb:144-153): every graph edge that crosses a task
boundary becomes Send/Recv, and ops that update a variable run *on the
variable's task* (Basics-Tutorial/Local-then-Global-Variables.ipynb:164-171).
Here each task owns a ``VariableHost``; an op on a remote variable is one
request executed by that host (``apply_sgd``, ``apply_adagrad``,
``acc_apply_grad`` ...), which gives the same "colocated with the variable"
semantics, including the deliberately unlocked read-modify-write of Hogwild
(Hogwild/Hogwild.py:1-8).

The host also carries the pieces TF keeps on the ps for synchronous training:
conditional accumulators and the shared token queue (SyncReplicasOptimizer,
Synchronous-SGD/ssgd.py:52-60), plus a key/value board used for rendezvous
(CUDA-IPC handle exchange on the GPU path) and done/heartbeat words.

Wire format: ``u32 header_len | JSON header | raw tensor bytes`` -- no pickle.
"""
from __future__ import annotations

import json
import socket
import socketserver
import struct
import threading
import time
from collections import deque
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

__all__ = ["VariableHost", "HostService", "HostClient", "RemoteError"]

_NP = {
    "float32": np.float32, "float64": np.float64, "float16": np.float16,
    "int32": np.int32, "int64": np.int64, "uint8": np.uint8, "bool": np.bool_,
    "int8": np.int8, "int16": np.int16,
}


class RemoteError(RuntimeError):
    pass


class AbortedError(RuntimeError):
    """Raised when a peer task went away mid-call (TF's ``AbortedError``, one of the two
    errors ``MonitoredSession`` recovers from, README.md:78)."""


class UnavailableError(AbortedError):
    """The peer task cannot be reached at all (TF's ``UnavailableError``: a ps that is not up
    yet or has been killed).  A subclass, so every ``except AbortedError`` recovery path
    handles both, as TF's ``_PREEMPTION_ERRORS`` tuple does."""


# --------------------------------------------------------------------------- #
# (de)serialisation helpers
# --------------------------------------------------------------------------- #
def _t2bytes(t: torch.Tensor) -> Tuple[dict, bytes]:
    t = t.detach()
    if t.is_cuda:
        t = t.cpu()
    if t.dtype == torch.bfloat16:
        meta = {"dtype": "bfloat16", "shape": list(t.shape)}
        return meta, t.contiguous().view(torch.int16).numpy().tobytes()
    a = t.contiguous().numpy()
    return {"dtype": str(a.dtype), "shape": list(a.shape)}, a.tobytes()


def _bytes2t(meta: dict, buf: bytes) -> torch.Tensor:
    if meta["dtype"] == "bfloat16":
        a = np.frombuffer(buf, dtype=np.int16).reshape(meta["shape"]).copy()
        return torch.from_numpy(a).view(torch.bfloat16)
    a = np.frombuffer(buf, dtype=_NP[meta["dtype"]]).reshape(meta["shape"]).copy()
    return torch.from_numpy(a)


def _recv_exact(sock: socket.socket, n: int) -> bytes:
    chunks, got = [], 0
    while got < n:
        b = sock.recv(min(n - got, 1 << 20))
        if not b:
            raise ConnectionError("peer closed")
        chunks.append(b)
        got += len(b)
    return b"".join(chunks)


def _send_msg(sock: socket.socket, header: dict, tensors: Sequence[torch.Tensor] = ()):
    metas, payloads = [], []
    for t in tensors:
        m, p = _t2bytes(t)
        m["nbytes"] = len(p)
        metas.append(m)
        payloads.append(p)
    header = dict(header)
    header["tensors"] = metas
    h = json.dumps(header).encode()
    sock.sendall(struct.pack("<I", len(h)) + h + b"".join(payloads))


def _recv_msg(sock: socket.socket) -> Tuple[dict, List[torch.Tensor]]:
    (hl,) = struct.unpack("<I", _recv_exact(sock, 4))
    header = json.loads(_recv_exact(sock, hl).decode())
    tensors = []
    for m in header.get("tensors", []):
        tensors.append(_bytes2t(m, _recv_exact(sock, m["nbytes"])))
    return header, tensors


# --------------------------------------------------------------------------- #
# Conditional accumulator + token queue (what SyncReplicas keeps on the ps)
# --------------------------------------------------------------------------- #
class _Accumulator:
    """[TF] ConditionalAccumulator: sums gradients whose ``local_step`` is not
    older than the accumulator's ``global_step``; ``take_grad(n)`` blocks until
    ``n`` were accepted, returns their *mean* and resets."""

    def __init__(self, shape, dtype):
        self.sum = torch.zeros(shape, dtype=dtype)
        self.count = 0
        self.global_step = 0
        self.dropped = 0
        self.cv = threading.Condition()

    def apply_grad(self, grad: torch.Tensor, local_step: int) -> bool:
        with self.cv:
            if local_step < self.global_step:
                self.dropped += 1
                return False          # stale: silently dropped
            self.sum += grad.to(self.sum.dtype)
            self.count += 1
            self.cv.notify_all()
            return True

    def take_grad(self, num_required: int, timeout: Optional[float]) -> Optional[torch.Tensor]:
        with self.cv:
            ok = self.cv.wait_for(lambda: self.count >= num_required, timeout=timeout)
            if not ok:
                return None
            out = self.sum / float(self.count)
            self.sum = torch.zeros_like(self.sum)
            self.count = 0
            self.global_step += 1
            return out

    def wait_ready(self, num_required: int, timeout: Optional[float]) -> bool:
        """Block until ``num_required`` gradients are in, WITHOUT taking them."""
        with self.cv:
            return self.cv.wait_for(lambda: self.count >= num_required, timeout=timeout)

    def set_global_step(self, step: int):
        with self.cv:
            self.global_step = int(step)


class _TokenQueue:
    """[TF] the shared FIFOQueue ``sync_token_q`` carrying ``global_step``."""

    def __init__(self):
        self.q = deque()
        self.cv = threading.Condition()
        self.closed = False

    def enqueue_many(self, values):
        with self.cv:
            self.q.extend(int(v) for v in values)
            self.cv.notify_all()

    def dequeue(self, timeout: Optional[float]):
        with self.cv:
            ok = self.cv.wait_for(lambda: self.q or self.closed, timeout=timeout)
            if not ok or (self.closed and not self.q):
                return None
            return self.q.popleft()

    def close(self):
        with self.cv:
            self.closed = True
            self.cv.notify_all()

    def size(self):
        with self.cv:
            return len(self.q)


# --------------------------------------------------------------------------- #
# VariableHost
# --------------------------------------------------------------------------- #
class VariableHost:
    """State owned by one task.  All mutating ops are *unlocked* with respect to
    each other on purpose (``use_locking=False`` is TF's default; Hogwild's
    races are the algorithm).  Only container bookkeeping takes ``_meta``.
    """

    def __init__(self, job_name: str = "local", task_index: int = 0):
        self.job_name, self.task_index = job_name, task_index
        self.vars: Dict[str, torch.Tensor] = {}
        self.accs: Dict[str, _Accumulator] = {}
        self.queues: Dict[str, _TokenQueue] = {}
        self.kv: Dict[str, str] = {}
        self.done: set = set()
        self.heartbeats: Dict[str, float] = {}
        self._meta = threading.Lock()
        self._kv_cv = threading.Condition()
        self.stats = {"push_bytes": 0, "pull_bytes": 0, "ops": 0}

    # -- variables ---------------------------------------------------------- #
    def assign(self, name: str, value: torch.Tensor) -> None:
        """Create-or-overwrite; marks the variable initialised."""
        v = value.detach().clone()
        with self._meta:
            cur = self.vars.get(name)
            if cur is not None and cur.shape == v.shape and cur.dtype == v.dtype:
                cur.copy_(v)
            else:
                self.vars[name] = v
        self.stats["push_bytes"] += v.numel() * v.element_size()

    def read(self, name: str) -> torch.Tensor:
        try:
            v = self.vars[name]
        except KeyError:
            raise KeyError("Attempting to use uninitialized value %s" % name)
        self.stats["pull_bytes"] += v.numel() * v.element_size()
        return v.clone()

    def read_many(self, names: Sequence[str]) -> List[torch.Tensor]:
        return [self.read(n) for n in names]

    def assign_many(self, names: Sequence[str], values: Sequence[torch.Tensor]) -> None:
        for n, v in zip(names, values):
            self.assign(n, v)

    def assign_add(self, name: str, delta: torch.Tensor) -> torch.Tensor:
        v = self.vars[name]
        v.add_(delta.to(v.dtype))
        return v.clone()

    def uninitialized(self, names: Sequence[str]) -> List[str]:
        return [n for n in names if n not in self.vars]

    def list_variables(self) -> List[str]:
        return sorted(self.vars)

    def delete(self, names: Sequence[str]) -> None:
        with self._meta:
            for n in names:
                self.vars.pop(n, None)

    # -- optimizer kernels executed on the variable's task (K3, K4) --------- #
    def apply_sgd(self, names: Sequence[str], grads: Sequence[torch.Tensor], lr: float) -> None:
        """``w -= lr*g`` for each variable, no lock (ApplyGradientDescent)."""
        for n, g in zip(names, grads):
            w = self.vars[n]
            w.sub_(g.to(w.dtype), alpha=lr)
            self.stats["push_bytes"] += g.numel() * g.element_size()
        self.stats["ops"] += 1

    def apply_adagrad(self, names: Sequence[str], accum_names: Sequence[str],
                      grads: Sequence[torch.Tensor], lr: float,
                      initial_accumulator_value: float = 0.1) -> None:
        """``acc += g*g; w -= lr*g/sqrt(acc)`` (ApplyAdagrad).  The slot is
        created on first use with TF's initial value 0.1."""
        for n, an, g in zip(names, accum_names, grads):
            w = self.vars[n]
            acc = self.vars.get(an)
            if acc is None:
                with self._meta:
                    acc = self.vars.setdefault(
                        an, torch.full_like(w, initial_accumulator_value))
            g = g.to(w.dtype)
            acc.addcmul_(g, g)
            w.addcdiv_(g, acc.sqrt(), value=-lr)
            self.stats["push_bytes"] += g.numel() * g.element_size()
        self.stats["ops"] += 1

    def counter_add(self, name: str, delta: int) -> int:
        """AssignAdd on an integer step variable; returns the new value."""
        v = self.vars[name]
        v.add_(int(delta))
        return int(v.item())

    # -- SyncReplicas state -------------------------------------------------- #
    def acc_create(self, name: str, shape, dtype: str = "float32") -> None:
        with self._meta:
            if name not in self.accs:
                self.accs[name] = _Accumulator(tuple(shape), getattr(torch, dtype))

    def acc_apply_grad(self, names: Sequence[str], grads: Sequence[torch.Tensor],
                       local_step: int) -> List[bool]:
        return [self.accs[n].apply_grad(g, local_step) for n, g in zip(names, grads)]

    def acc_take_grad(self, names: Sequence[str], num_required: int,
                      timeout: Optional[float] = None) -> Optional[List[torch.Tensor]]:
        """All-or-nothing take over this host's accumulators.

        Phase 1 waits (within ``timeout`` overall) until EVERY accumulator holds
        ``num_required`` gradients and drains none; only then phase 2 takes them all.  A
        time-out therefore never leaves some accumulators reset (sum cleared, global_step
        bumped) and others not -- that state drops the workers' next pushes as stale and
        starves the token queue.  Counts only grow between the phases: the chief's
        aggregation thread is the only taker.
        """
        deadline = None if timeout is None else time.time() + timeout
        for n in names:
            left = None if deadline is None else max(0.0, deadline - time.time())
            if not self.accs[n].wait_ready(num_required, left):
                return None
        out = []
        for n in names:
            g = self.accs[n].take_grad(num_required, None)
            out.append(g)
        return out

    def acc_set_global_step(self, names: Sequence[str], step: int) -> None:
        for n in names:
            self.accs[n].set_global_step(step)

    def acc_info(self, name: str) -> dict:
        a = self.accs[name]
        return {"count": a.count, "global_step": a.global_step, "dropped": a.dropped}

    def queue_create(self, name: str) -> None:
        with self._meta:
            self.queues.setdefault(name, _TokenQueue())

    def queue_enqueue_many(self, name: str, values: Sequence[int]) -> None:
        self.queue_create(name)
        self.queues[name].enqueue_many(values)

    def queue_dequeue(self, name: str, timeout: Optional[float] = None) -> Optional[int]:
        self.queue_create(name)
        return self.queues[name].dequeue(timeout)

    def queue_size(self, name: str) -> int:
        self.queue_create(name)
        return self.queues[name].size()

    def queue_close(self, name: str) -> None:
        self.queue_create(name)
        self.queues[name].close()

    # -- rendezvous board / liveness ----------------------------------------- #
    def kv_set(self, key: str, value: str) -> None:
        with self._kv_cv:
            self.kv[key] = value
            self._kv_cv.notify_all()

    def kv_get(self, key: str, timeout: Optional[float] = None) -> Optional[str]:
        with self._kv_cv:
            ok = self._kv_cv.wait_for(lambda: key in self.kv, timeout=timeout)
            return self.kv[key] if ok else None

    def kv_add(self, key: str, delta: int) -> int:
        with self._kv_cv:
            v = int(self.kv.get(key, "0")) + int(delta)
            self.kv[key] = str(v)
            self._kv_cv.notify_all()
            return v

    def mark_done(self, who: str) -> int:
        with self._meta:
            self.done.add(who)
            return len(self.done)

    def num_done(self) -> int:
        return len(self.done)

    def heartbeat(self, who: str) -> None:
        self.heartbeats[who] = time.time()

    def stale_tasks(self, max_age: float) -> List[str]:
        now = time.time()
        return sorted(k for k, t in self.heartbeats.items() if now - t > max_age)

    def get_stats(self) -> dict:
        return dict(self.stats)

    def ping(self) -> str:
        return "%s:%d" % (self.job_name, self.task_index)


# Methods a client may invoke.  (name, blocking?)  -- nothing outside this table
# is reachable from the wire.
_EXPORTED = {
    "assign", "read", "read_many", "assign_many", "assign_add", "uninitialized",
    "list_variables", "delete", "apply_sgd", "apply_adagrad", "counter_add",
    "acc_create", "acc_apply_grad", "acc_take_grad", "acc_set_global_step", "acc_info",
    "queue_create", "queue_enqueue_many", "queue_dequeue", "queue_size", "queue_close",
    "kv_set", "kv_get", "kv_add", "mark_done", "num_done", "heartbeat", "stale_tasks",
    "get_stats", "ping",
}


def _split_args(args):
    """Replace tensors (and flat lists of tensors) in ``args`` by placeholders
    and return them separately so they travel as raw bytes."""
    tensors, enc = [], []
    for a in args:
        if isinstance(a, torch.Tensor):
            enc.append({"__t__": len(tensors)})
            tensors.append(a)
        elif isinstance(a, (list, tuple)) and a and all(isinstance(x, torch.Tensor) for x in a):
            enc.append({"__tl__": [len(tensors) + i for i in range(len(a))]})
            tensors.extend(a)
        else:
            enc.append(a)
    return enc, tensors


def _join_args(enc, tensors):
    out = []
    for a in enc:
        if isinstance(a, dict) and "__t__" in a:
            out.append(tensors[a["__t__"]])
        elif isinstance(a, dict) and "__tl__" in a:
            out.append([tensors[i] for i in a["__tl__"]])
        else:
            out.append(a)
    return out


class _Handler(socketserver.BaseRequestHandler):
    def handle(self):
        host: VariableHost = self.server.host  # type: ignore[attr-defined]
        sock = self.request
        sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        while True:
            try:
                header, tensors = _recv_msg(sock)
            except (ConnectionError, OSError, struct.error):
                return
            op = header.get("op")
            try:
                if op not in _EXPORTED:
                    raise RemoteError("unknown op %r" % (op,))
                res = getattr(host, op)(*_join_args(header.get("args", []), tensors))
                enc, out_t = _split_args([res])
                _send_msg(sock, {"ok": True, "result": enc[0]}, out_t)
            except Exception as e:  # report to the caller, keep serving
                try:
                    _send_msg(sock, {"ok": False, "error": "%s: %s" % (type(e).__name__, e)})
                except OSError:
                    return


class _TCPServer(socketserver.ThreadingTCPServer):
    allow_reuse_address = True
    daemon_threads = True
    request_queue_size = 128


class HostService:
    """Serves a ``VariableHost`` on ``host:port`` from a daemon thread."""

    def __init__(self, host: VariableHost, addr: str, port: int):
        bind = "127.0.0.1" if addr in ("localhost", "127.0.0.1") else addr
        self._srv = _TCPServer((bind, port), _Handler)
        self._srv.host = host  # type: ignore[attr-defined]
        self.port = self._srv.server_address[1]
        self._thread = threading.Thread(target=self._srv.serve_forever,
                                        kwargs={"poll_interval": 0.05}, daemon=True,
                                        name="dtg-host-%d" % self.port)

    def start(self):
        self._thread.start()

    def stop(self):
        try:
            self._srv.shutdown()
            self._srv.server_close()
        except Exception:
            pass


class HostClient:
    """RPC proxy with the ``VariableHost`` method surface.  One connection per
    calling thread, so a queue-runner thread blocked in ``acc_take_grad`` does
    not stall the training thread."""

    def __init__(self, addr: str, port: int, connect_timeout: float = 60.0):
        self._addr = "127.0.0.1" if addr == "localhost" else addr
        self._port = port
        self._connect_timeout = connect_timeout
        self._tls = threading.local()
        self._all: List[socket.socket] = []
        self._lock = threading.Lock()

    def _sock(self) -> socket.socket:
        s = getattr(self._tls, "sock", None)
        if s is not None:
            return s
        t0 = time.time()
        while True:
            try:
                s = socket.create_connection((self._addr, self._port), timeout=5.0)
                break
            except OSError:
                if time.time() - t0 > self._connect_timeout:
                    raise UnavailableError("cannot reach task at %s:%d" % (self._addr, self._port))
                time.sleep(0.05)
        s.settimeout(None)
        s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        self._tls.sock = s
        with self._lock:
            self._all.append(s)
        return s

    def _call(self, op: str, *args):
        enc, tensors = _split_args(args)
        try:
            s = self._sock()
            _send_msg(s, {"op": op, "args": enc}, tensors)
            header, out_t = _recv_msg(s)
        except (ConnectionError, OSError) as e:
            self._tls.sock = None
            raise AbortedError("task at %s:%d went away during %s (%s)" % (
                self._addr, self._port, op, e))
        if not header.get("ok"):
            err = header.get("error", "")
            if err.startswith("KeyError"):
                raise KeyError(err)
            raise RemoteError(err)
        return _join_args([header["result"]], out_t)[0]

    def __getattr__(self, name):
        if name in _EXPORTED:
            return lambda *a: self._call(name, *a)
        raise AttributeError(name)

    def close(self):
        with self._lock:
            for s in self._all:
                try:
                    s.close()
                except OSError:
                    pass
            self._all.clear()

"""Cluster description, device names and the per-task ``Server``.

Capabilities mirrored (reference cites are /root/reference paths):

* ``ClusterSpec({'ps': [...], 'worker': [...]})`` -- Hogwild/Hogwild.py:20-23,
  DOWNPOUR/DOWNPOUR.py:24-28 (the dict is also used to count ps / workers).
* ``Server(cluster, job_name, task_index, config)``, ``server.join()``,
  ``server.target`` -- Hogwild/Hogwild.py:25-33,50.
* ``Server.create_local_server()`` and device listing --
  Basics-Tutorial/Servers.ipynb:198,246-253.
* ``ConfigProto`` / ``GPUOptions`` --
  Multiple-GPUs-Single-Machine/dist_mult_gpu_sing_mach.py:45-49.

Design: one OS process per task.  Every task runs a small TCP *variable host*
(see ``parallel/rpc.py``) on the address the cluster spec gives it.  On CPU the
host is the transport itself (the star topology of the reference: workers talk
to the ps, never to each other).  On B200 the same service is only the control
plane -- rendezvous, CUDA-IPC handle exchange, done/heartbeat signalling --
while the data plane is NVLink peer memory touched directly by worker kernels
(``parallel/fabric.py``).
"""
from __future__ import annotations

import dataclasses
import re
import socket
import threading
import time
from typing import Dict, List, Optional, Union

__all__ = [
    "ClusterSpec", "Server", "ConfigProto", "GPUOptions", "DeviceSpec",
    "DeviceAttributes", "canonical_device",
]


# --------------------------------------------------------------------------- #
# Device specs
# --------------------------------------------------------------------------- #
_DEV_RE = re.compile(
    r"^(?:/job:(?P<job>[A-Za-z_][\w]*))?"
    r"(?:/replica:(?P<replica>\d+))?"
    r"(?:/task:(?P<task>\d+))?"
    r"(?:/(?:device:)?(?P<dtype>[A-Za-z]+):(?P<dindex>\d+|\*))?$"
)


@dataclasses.dataclass
class DeviceSpec:
    """Partially specified device, e.g. ``/job:ps/task:0`` or ``/gpu:0``.

    ``merge_from`` implements the nested ``tf.device`` rule: the inner scope's
    fields override the outer scope's, unspecified fields are inherited
    (Basics-Tutorial/Local-then-Global-Variables.ipynb:104,117,127,133).
    """

    job: Optional[str] = None
    replica: Optional[int] = None
    task: Optional[int] = None
    device_type: Optional[str] = None
    device_index: Optional[int] = None

    @classmethod
    def from_string(cls, spec: Optional[str]) -> "DeviceSpec":
        if not spec:
            return cls()
        m = _DEV_RE.match(spec.strip())
        if m is None:
            raise ValueError("malformed device spec %r" % (spec,))
        d = m.groupdict()
        dindex = d["dindex"]
        return cls(
            job=d["job"],
            replica=int(d["replica"]) if d["replica"] is not None else None,
            task=int(d["task"]) if d["task"] is not None else None,
            device_type=d["dtype"].upper() if d["dtype"] else None,
            device_index=int(dindex) if dindex not in (None, "*") else None,
        )

    def merge_from(self, inner: "DeviceSpec") -> "DeviceSpec":
        out = dataclasses.replace(self)
        for f in ("job", "replica", "task", "device_type", "device_index"):
            v = getattr(inner, f)
            if v is not None:
                setattr(out, f, v)
        return out

    def to_string(self) -> str:
        s = ""
        if self.job is not None:
            s += "/job:%s" % self.job
        if self.replica is not None:
            s += "/replica:%d" % self.replica
        if self.task is not None:
            s += "/task:%d" % self.task
        if self.device_type is not None:
            s += "/device:%s:%s" % (
                self.device_type,
                "*" if self.device_index is None else str(self.device_index))
        return s

    __str__ = to_string


def canonical_device(spec: Union[str, DeviceSpec, None]) -> str:
    """Fully qualified form used when printing placements
    (``/job:ps/replica:0/task:0/device:CPU:0``)."""
    d = spec if isinstance(spec, DeviceSpec) else DeviceSpec.from_string(spec)
    d = dataclasses.replace(d)
    if d.replica is None:
        d.replica = 0
    if d.task is None:
        d.task = 0
    if d.device_type is None:
        d.device_type = "CPU"
    if d.device_index is None:
        d.device_index = 0
    return d.to_string()


@dataclasses.dataclass
class DeviceAttributes:
    """One row of ``Session.list_devices()``
    (Basics-Tutorial/Servers.ipynb:246: ``_DeviceAttributes(name, type, mem)``)."""
    name: str
    device_type: str
    memory_limit_bytes: int

    def __repr__(self):
        return "_DeviceAttributes(%s, %s, %d)" % (
            self.name, self.device_type, self.memory_limit_bytes)


# --------------------------------------------------------------------------- #
# Config
# --------------------------------------------------------------------------- #
@dataclasses.dataclass
class GPUOptions:
    """Multiple-GPUs-Single-Machine/dist_mult_gpu_sing_mach.py:45-47.

    ``visible_device_list`` picks the physical GPU(s) of this task (the
    reference pins worker *k* to GPU *k*).  ``allow_growth`` / ``allocator_type``
    are accepted for source compatibility; torch's caching allocator always
    grows on demand.
    """
    allow_growth: bool = True
    allocator_type: str = "BFC"
    visible_device_list: str = ""
    per_process_gpu_memory_fraction: float = 1.0


@dataclasses.dataclass
class ConfigProto:
    """Runtime options (DOWNPOUR/DOWNPOUR.py:21)."""
    log_device_placement: bool = False
    allow_soft_placement: bool = False
    gpu_options: GPUOptions = dataclasses.field(default_factory=GPUOptions)
    #: seconds a blocking control-plane call may wait before raising
    operation_timeout_in_ms: int = 0


# --------------------------------------------------------------------------- #
# ClusterSpec
# --------------------------------------------------------------------------- #
class ClusterSpec:
    """Job name -> ordered list of ``host:port``; the task index is the list
    position (Hogwild/Hogwild.py:20-23).

    Accepts a dict, another ``ClusterSpec`` or the CLI short form
    ``"ps=1,worker=2"`` (ports are then assigned from ``base_port`` upwards in
    the reference's order: ps first, i.e. 2222, 2223, 2224 ...).
    """

    def __init__(self, cluster: Union[dict, "ClusterSpec", str],
                 base_port: int = 2222, host: str = "localhost"):
        if isinstance(cluster, ClusterSpec):
            cluster = cluster.as_dict()
        if isinstance(cluster, str):
            cluster = self._parse_short(cluster, base_port, host)
        if not isinstance(cluster, dict):
            raise TypeError("ClusterSpec expects a dict, a ClusterSpec or 'ps=N,worker=M'")
        self._spec: Dict[str, Dict[int, str]] = {}
        for job, tasks in cluster.items():
            if isinstance(tasks, dict):
                self._spec[job] = {int(i): str(a) for i, a in tasks.items()}
            elif isinstance(tasks, (list, tuple)):
                self._spec[job] = {i: str(a) for i, a in enumerate(tasks)}
            else:
                raise TypeError("tasks of job %r must be a list or a dict" % job)

    @staticmethod
    def _parse_short(s: str, base_port: int, host: str) -> dict:
        counts = {}
        for part in s.split(","):
            k, v = part.split("=")
            counts[k.strip()] = int(v)
        out, port = {}, base_port
        for job in sorted(counts, key=lambda j: (j != "ps", j)):
            out[job] = ["%s:%d" % (host, port + i) for i in range(counts[job])]
            port += counts[job]
        return out

    # -- queries ------------------------------------------------------------ #
    @property
    def jobs(self) -> List[str]:
        return list(self._spec.keys())

    def num_tasks(self, job_name: str) -> int:
        return len(self._job(job_name))

    def task_indices(self, job_name: str) -> List[int]:
        return sorted(self._job(job_name))

    def task_address(self, job_name: str, task_index: int) -> str:
        job = self._job(job_name)
        if task_index not in job:
            raise ValueError("no task %d in job %r" % (task_index, job_name))
        return job[task_index]

    def job_tasks(self, job_name: str) -> List[str]:
        job = self._job(job_name)
        return [job[i] for i in sorted(job)]

    def as_dict(self) -> dict:
        out = {}
        for job, tasks in self._spec.items():
            idx = sorted(tasks)
            if idx == list(range(len(idx))):
                out[job] = [tasks[i] for i in idx]
            else:
                out[job] = dict(tasks)
        return out

    def as_cluster_def(self) -> str:
        """Text form shown by ``server.server_def``
        (Basics-Tutorial/Servers.ipynb:144-153)."""
        lines = ["cluster {"]
        for job in self._spec:
            lines.append("  job {")
            lines.append('    name: "%s"' % job)
            for i in sorted(self._spec[job]):
                lines.append("    tasks {")
                lines.append("      key: %d" % i)
                lines.append('      value: "%s"' % self._spec[job][i])
                lines.append("    }")
            lines.append("  }")
        lines.append("}")
        return "\n".join(lines)

    def all_tasks(self):
        for job in self._spec:
            for i in sorted(self._spec[job]):
                yield job, i, self._spec[job][i]

    def global_rank(self, job_name: str, task_index: int) -> int:
        """Dense rank over all tasks, ps first -- the role -> GPU map on one
        8xB200 host is ``rank % n_gpus`` unless ``visible_device_list`` says
        otherwise."""
        for r, (j, i, _) in enumerate(sorted(
                self.all_tasks(), key=lambda t: (t[0] != "ps", t[0], t[1]))):
            if j == job_name and i == task_index:
                return r
        raise ValueError("unknown task %s:%d" % (job_name, task_index))

    def _job(self, job_name):
        if job_name not in self._spec:
            raise ValueError("no job %r in cluster (jobs: %s)" % (job_name, self.jobs))
        return self._spec[job_name]

    def __eq__(self, other):
        return isinstance(other, ClusterSpec) and self._spec == other._spec

    def __bool__(self):
        return bool(self._spec)

    def __repr__(self):
        return "ClusterSpec(%r)" % (self.as_dict(),)


# --------------------------------------------------------------------------- #
# Server
# --------------------------------------------------------------------------- #
SCHEME = "grpc"   # target strings keep the reference's look: 'grpc://localhost:2222'


def _split_addr(addr: str):
    host, port = addr.rsplit(":", 1)
    return host, int(port)


class Server:
    """In-process task server (Hogwild/Hogwild.py:25-33).

    Starting a ``Server`` binds this task's address, starts the variable-host
    service thread and registers the task as the process's *runtime*, so that
    variables placed on this task are stored in-process and variables placed on
    other tasks are reached through their hosts.

    ``join()`` blocks the ps role.  The reference's ps never exits
    (README.md:55-59, killed with ``pkill``); here ``join()`` returns once every
    worker has signalled ``done`` unless ``linger=True``.
    """

    def __init__(self, cluster: Union[ClusterSpec, dict], job_name: str = "worker",
                 task_index: int = 0, config: Optional[ConfigProto] = None,
                 start: bool = True, protocol: str = "grpc"):
        from .parallel import rpc  # local import: rpc imports variables

        self.cluster = cluster if isinstance(cluster, ClusterSpec) else ClusterSpec(cluster)
        self.job_name = job_name
        self.task_index = int(task_index)
        self.config = config or ConfigProto()
        self.protocol = protocol
        self._address = self.cluster.task_address(job_name, self.task_index)
        self._host = rpc.VariableHost(job_name, self.task_index)
        self._service: Optional["rpc.HostService"] = None
        self._started = False
        self._clients: Dict[str, "rpc.HostClient"] = {}
        self._lock = threading.Lock()
        self._gpu_index = self._pick_gpu()
        if start:
            self.start()

    # -- lifecycle ---------------------------------------------------------- #
    def start(self):
        from .parallel import rpc
        if self._started:
            return
        host, port = _split_addr(self._address)
        self._service = rpc.HostService(self._host, host, port)
        self._service.start()
        if port == 0:  # ephemeral (create_local_server)
            self._address = "%s:%d" % (host, self._service.port)
            self.cluster._spec[self.job_name][self.task_index] = self._address
        self._started = True
        from . import runtime
        runtime.set_current_server(self)

    def join(self, linger: bool = False, poll_secs: float = 0.05,
             timeout: Optional[float] = None):
        """Block while passively hosting variables (Hogwild/Hogwild.py:28)."""
        if not self._started:
            self.start()
        n_workers = self.cluster.num_tasks("worker") if "worker" in self.cluster.jobs else 0
        t0 = time.time()
        while True:
            if not linger and n_workers and self._host.num_done() >= n_workers:
                break
            if timeout is not None and time.time() - t0 > timeout:
                break
            time.sleep(poll_secs)
        self.stop()

    def stop(self):
        if self._service is not None:
            self._service.stop()
            self._service = None
        for c in self._clients.values():
            c.close()
        self._clients.clear()
        self._started = False
        from . import runtime
        if runtime.current_server(required=False) is self:
            runtime.set_current_server(None)

    # -- properties --------------------------------------------------------- #
    @property
    def target(self) -> str:
        return "%s://%s" % (self.protocol, self._address)

    @property
    def server_def(self) -> str:
        return "%s\njob_name: \"%s\"\ntask_index: %d\nprotocol: \"%s\"" % (
            self.cluster.as_cluster_def(), self.job_name, self.task_index, self.protocol)

    @property
    def host(self):
        return self._host

    @property
    def gpu_index(self) -> Optional[int]:
        return self._gpu_index

    @classmethod
    def create_local_server(cls, config: Optional[ConfigProto] = None,
                            start: bool = True) -> "Server":
        """Single-process cluster on an ephemeral port
        (Basics-Tutorial/Servers.ipynb:198,219)."""
        return cls({"local": ["localhost:0"]}, job_name="local", task_index=0,
                   config=config, start=start)

    # -- devices ------------------------------------------------------------ #
    def _pick_gpu(self) -> Optional[int]:
        try:
            import torch
            if not torch.cuda.is_available():
                return None
            n = torch.cuda.device_count()
        except Exception:  # pragma: no cover
            return None
        vis = self.config.gpu_options.visible_device_list
        if vis:
            first = vis.split(",")[0].strip()
            if first == "-1":
                return None
            return int(first) % n
        return self.cluster.global_rank(self.job_name, self.task_index) % n

    def list_devices(self) -> List[DeviceAttributes]:
        """Devices of every task in the cluster, fully qualified
        (Basics-Tutorial/Servers.ipynb:251-253)."""
        out = []
        try:
            import psutil
            cpu_mem = int(psutil.virtual_memory().total)
        except Exception:  # pragma: no cover
            cpu_mem = 268435456
        for job, idx, _ in self.cluster.all_tasks():
            out.append(DeviceAttributes(
                "/job:%s/replica:0/task:%d/device:CPU:0" % (job, idx), "CPU", cpu_mem))
        if self._gpu_index is not None:
            import torch
            props = torch.cuda.get_device_properties(self._gpu_index)
            out.append(DeviceAttributes(
                "/job:%s/replica:0/task:%d/device:GPU:0" % (self.job_name, self.task_index),
                "GPU", int(props.total_memory)))
        return out

    # -- routing ------------------------------------------------------------ #
    def is_local(self, job: Optional[str], task: Optional[int]) -> bool:
        return (job is None or job == self.job_name) and \
               (task is None or task == self.task_index)

    def client_for(self, job: str, task: int):
        """Control/data-plane client for another task's variable host."""
        from .parallel import rpc
        key = "%s:%d" % (job, task)
        with self._lock:
            c = self._clients.get(key)
            if c is None:
                host, port = _split_addr(self.cluster.task_address(job, task))
                c = rpc.HostClient(host, port)
                self._clients[key] = c
            return c

    def host_for(self, job: Optional[str], task: Optional[int]):
        """The object that executes variable ops for ``/job:J/task:T``: the
        in-process host when the device is this task, else an RPC proxy."""
        if self.is_local(job, task):
            return self._host
        return self.client_for(job, task if task is not None else 0)

    def __repr__(self):
        return "<Server %s job=%s task=%d>" % (self.target, self.job_name, self.task_index)


def wait_for_port(host: str, port: int, timeout: float = 30.0) -> bool:
    """Poll until something listens on ``host:port`` (the notebooks check this
    with ``lsof``, Basics-Tutorial/Servers.ipynb:112-120)."""
    t0 = time.time()
    while time.time() - t0 < timeout:
        try:
            with socket.create_connection((host, port), timeout=0.25):
                return True
        except OSError:
            time.sleep(0.05)
    return False

"""Process-wide runtime: which ``Server`` this process is.

In the reference a worker process builds its graph and then opens a session on
``server.target`` -- its own in-process master (Hogwild/Hogwild.py:47-50).  The
equivalent here is a process-global "current server"; variable ops route through
it (local host or RPC proxy).  If no server was started (C1,
Non-Distributed_Setup.py) an implicit single-task local runtime is used.
"""
from __future__ import annotations

import threading

_lock = threading.Lock()
_current = None
_implicit = None


class _ImplicitLocal:
    """Stand-in for a Server when the program is not distributed at all."""

    job_name = "localhost"
    task_index = 0
    gpu_index = None
    cluster = None
    config = None

    def __init__(self):
        from .parallel.rpc import VariableHost
        self.host = VariableHost("localhost", 0)
        try:
            import torch
            if torch.cuda.is_available():
                self.gpu_index = 0
        except Exception:  # pragma: no cover
            pass

    target = ""

    def is_local(self, job, task):
        return True

    def host_for(self, job, task):
        return self.host

    def list_devices(self):
        from .cluster import DeviceAttributes
        return [DeviceAttributes("/job:localhost/replica:0/task:0/device:CPU:0", "CPU", 268435456)]


def set_current_server(server) -> None:
    global _current
    with _lock:
        _current = server


def current_server(required: bool = True):
    global _implicit
    with _lock:
        if _current is not None:
            return _current
        if not required:
            return None
        if _implicit is None:
            _implicit = _ImplicitLocal()
        return _implicit


def reset() -> None:
    """Forget the implicit local runtime (tests)."""
    global _implicit, _current
    with _lock:
        _implicit = None
        _current = None

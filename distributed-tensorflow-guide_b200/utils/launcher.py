"""Start every role of a cluster from one command -- what the reference's
``run.sh`` files do with three backgrounded ``python`` lines
(Hogwild/run.sh:2-4) and what ``dist_mult_gpu_sing_mach.sh:2-7`` does with
``CUDA_VISIBLE_DEVICES`` pinning (ps hidden from GPUs, worker k -> GPU k)."""
from __future__ import annotations

import os
import subprocess
import sys
import time
from typing import Dict, List, Optional, Sequence

from ..cluster import ClusterSpec

__all__ = ["launch_cluster", "wait_all"]


def launch_cluster(script_argv: Sequence[str], cluster: ClusterSpec, pin_gpus: bool = False,
                   env: Optional[Dict[str, str]] = None, stdout=None,
                   python: str = sys.executable) -> List[subprocess.Popen]:
    """``script_argv`` is e.g. ``['examples/hogwild.py']`` or ``['-m', 'dtg_b200', ...]``;
    ``--job_name J --task_index T`` is appended per role.  ps first, like run.sh."""
    procs = []
    order = sorted(cluster.all_tasks(), key=lambda t: (t[0] != "ps", t[0], t[1]))
    n_gpu_workers = 0
    for job, idx, _ in order:
        e = dict(os.environ)
        e.update(env or {})
        if pin_gpus:
            if job == "ps":
                e["CUDA_VISIBLE_DEVICES"] = "-1"
            else:
                e["CUDA_VISIBLE_DEVICES"] = str(n_gpu_workers)
                n_gpu_workers += 1
        cmd = [python] + list(script_argv) + ["--job_name", job, "--task_index", str(idx)]
        procs.append(subprocess.Popen(cmd, env=e, stdout=stdout, stderr=subprocess.STDOUT))
        if job == "ps":
            time.sleep(0.05)
    return procs


def wait_all(procs: Sequence[subprocess.Popen], timeout: float = 600.0) -> List[int]:
    deadline = time.time() + timeout
    codes = []
    for p in procs:
        try:
            codes.append(p.wait(timeout=max(0.1, deadline - time.time())))
        except subprocess.TimeoutExpired:
            p.kill()
            codes.append(-9)
    return codes

"""Build ``libdtg_b200.so`` in-tree with nvcc for sm_100a.

No JIT cache, no torch C++ headers: every kernel is reached through a C ABI
(``include/dtg_b200.h``) and loaded with ctypes, so a build is a handful of
``nvcc -c`` calls (seconds each) plus one link.  The ``.so`` is git-ignored but
travels to the GPU box with the snapshot.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from typing import List

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdtg_b200.so")
OBJ_DIR = os.path.join(HERE, "_build")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def sources() -> List[str]:
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest(paths: List[str]) -> str:
    """Content hash of the inputs of one object (sources, headers, flags): mtimes do not survive
    a snapshot copy to another box, contents do."""
    h = hashlib.sha256(" ".join(NVCC_FLAGS).encode())
    for p in sorted(paths):
        h.update(os.path.basename(p).encode())
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _stale(target: str, deps: List[str]) -> bool:
    """True unless `target` exists and was built from exactly these inputs."""
    stamp = target + ".sha256"
    if not os.path.exists(target) or not os.path.exists(stamp):
        return True
    with open(stamp) as f:
        return f.read().strip() != _digest(deps)


def _stamp(target: str, deps: List[str]) -> None:
    with open(target + ".sha256", "w") as f:
        f.write(_digest(deps))


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ_DIR, exist_ok=True)
    nvcc = _nvcc()
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    srcs = sources()
    objs = [os.path.join(OBJ_DIR, os.path.basename(s)[:-3] + ".o") for s in srcs]

    def compile_one(pair):
        src, obj = pair
        if not force and not _stale(obj, [src] + headers):
            return ""
        cmd = [nvcc] + NVCC_FLAGS + ["-c", src, "-o", obj]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (src, r.stdout))
        with open(obj[:-2] + ".ptxas.log", "w") as f:
            f.write(r.stdout)
        _stamp(obj, [src] + headers)
        return r.stdout

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        logs = list(ex.map(compile_one, zip(srcs, objs)))
    if verbose:
        for l in logs:
            if l:
                print(l)
    all_inputs = srcs + headers
    if force or any(logs) or _stale(LIB, all_inputs):
        cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s" % r.stdout)
        _stamp(LIB, all_inputs)
    return LIB


def is_current() -> bool:
    """Does the shipped library match the sources next to it?  (ops.lib() refuses a stale one.)"""
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    return not _stale(LIB, sources() + headers)


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))

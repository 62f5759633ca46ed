"""Checkpoints keyed by TF variable name.

Reference behaviour ([TF] Saver, invoked through
``MonitoredTrainingSession(checkpoint_dir='logdir', save_checkpoint_secs=1)``
DOWNPOUR/DOWNPOUR.py:121-127 and ``Supervisor(logdir=..., save_model_secs=30)``
Hogwild/Hogwild.py:47-49): only the chief saves, only GLOBAL variables (which
includes ``global_step`` and ps-side optimizer slots such as
``g/Variable/Adagrad``), files are ``model.ckpt-<step>.*`` plus a ``checkpoint``
text file naming the latest, and a new session restores the latest checkpoint
automatically.

Kept: directory layout, the ``checkpoint`` state file (same text syntax as TF),
name keys, ``max_to_keep`` rotation.  Payload is our own single-file container
(``<prefix>.dtg``): ``u64 header_len | JSON {name: {dtype, shape, offset,
nbytes}} | raw little-endian bytes`` -- mmap-friendly and pickle-free.
"""
from __future__ import annotations

import json
import os
import re
import struct
import time
from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np
import torch

from .. import framework as fw
from ..framework import GraphKeys, Variable

__all__ = ["Saver", "latest_checkpoint", "get_checkpoint_state", "load_checkpoint",
           "save_tensors", "checkpoint_exists", "SUFFIX"]

SUFFIX = ".dtg"
_MAGIC = b"DTGCKPT1"

_TORCH2NP = {torch.float32: "float32", torch.float64: "float64", torch.float16: "float16",
             torch.int32: "int32", torch.int64: "int64", torch.uint8: "uint8",
             torch.bool: "bool", torch.int8: "int8", torch.int16: "int16"}


def save_tensors(path: str, tensors: Dict[str, torch.Tensor]) -> None:
    """Write ``{name: tensor}`` atomically to ``path``."""
    header, blobs, off = {}, [], 0
    for name in sorted(tensors):
        t = tensors[name].detach()
        if t.is_cuda:
            t = t.cpu()
        t = t.contiguous()
        if t.dtype == torch.bfloat16:
            dt, raw = "bfloat16", t.view(torch.int16).numpy().tobytes()
        else:
            dt, raw = _TORCH2NP[t.dtype], t.numpy().tobytes()
        header[name] = {"dtype": dt, "shape": list(t.shape), "offset": off, "nbytes": len(raw)}
        blobs.append(raw)
        off += len(raw)
    h = json.dumps(header, sort_keys=True).encode()
    tmp = path + ".tmp.%d" % os.getpid()
    with open(tmp, "wb") as f:
        f.write(_MAGIC)
        f.write(struct.pack("<Q", len(h)))
        f.write(h)
        for b in blobs:
            f.write(b)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)


def load_checkpoint(path: str, names: Optional[Iterable[str]] = None) -> Dict[str, torch.Tensor]:
    if not path.endswith(SUFFIX) and not os.path.exists(path + SUFFIX):
        from . import tf_bundle
        if tf_bundle.is_bundle(path):          # a TF Saver V2 checkpoint (model.ckpt-N.index/.data)
            return tf_bundle.read_bundle(path, names)
    if not path.endswith(SUFFIX):
        path = path + SUFFIX
    with open(path, "rb") as f:
        if f.read(len(_MAGIC)) != _MAGIC:
            raise ValueError("%s is not a dtg checkpoint" % path)
        (hl,) = struct.unpack("<Q", f.read(8))
        header = json.loads(f.read(hl).decode())
        base = f.tell()
        want = set(names) if names is not None else None
        out = {}
        for name, m in header.items():
            if want is not None and name not in want:
                continue
            f.seek(base + m["offset"])
            raw = f.read(m["nbytes"])
            if m["dtype"] == "bfloat16":
                a = np.frombuffer(raw, dtype=np.int16).reshape(m["shape"]).copy()
                out[name] = torch.from_numpy(a).view(torch.bfloat16)
            else:
                a = np.frombuffer(raw, dtype=np.dtype(m["dtype"])).reshape(m["shape"]).copy()
                out[name] = torch.from_numpy(a)
    return out


def list_checkpoint_variables(path: str) -> Dict[str, tuple]:
    if not path.endswith(SUFFIX):
        path = path + SUFFIX
    with open(path, "rb") as f:
        f.read(len(_MAGIC))
        (hl,) = struct.unpack("<Q", f.read(8))
        header = json.loads(f.read(hl).decode())
    return {k: (v["dtype"], tuple(v["shape"])) for k, v in header.items()}


# ---- 'checkpoint' state file (TF text syntax) -------------------------------- #
def _state_path(d: str, latest_filename: Optional[str] = None) -> str:
    return os.path.join(d, latest_filename or "checkpoint")


def update_checkpoint_state(d: str, model_checkpoint_path: str,
                            all_paths: Sequence[str], latest_filename=None) -> None:
    lines = ['model_checkpoint_path: "%s"' % os.path.basename(model_checkpoint_path)]
    for p in all_paths:
        lines.append('all_model_checkpoint_paths: "%s"' % os.path.basename(p))
    tmp = _state_path(d, latest_filename) + ".tmp.%d" % os.getpid()
    with open(tmp, "w") as f:
        f.write("\n".join(lines) + "\n")
    os.replace(tmp, _state_path(d, latest_filename))


def get_checkpoint_state(d: str, latest_filename=None) -> Optional[dict]:
    p = _state_path(d, latest_filename)
    if not os.path.exists(p):
        return None
    state = {"model_checkpoint_path": None, "all_model_checkpoint_paths": []}
    with open(p) as f:
        for line in f:
            m = re.match(r'\s*(\w+):\s*"(.*)"\s*$', line)
            if not m:
                continue
            k, v = m.group(1), m.group(2)
            v = v if os.path.isabs(v) else os.path.join(d, v)
            if k == "model_checkpoint_path":
                state[k] = v
            elif k == "all_model_checkpoint_paths":
                state[k].append(v)
    return state


def checkpoint_exists(prefix: str) -> bool:
    return os.path.exists(prefix + SUFFIX) or os.path.exists(prefix + ".index")


def latest_checkpoint(d: str, latest_filename=None) -> Optional[str]:
    st = get_checkpoint_state(d, latest_filename)
    if st and st["model_checkpoint_path"] and checkpoint_exists(st["model_checkpoint_path"]):
        return st["model_checkpoint_path"]
    return None


class Saver:
    """Save / restore GLOBAL variables by name.

    ``var_list``: a sequence of variables (saved under ``v.op.name``) or, as in TF, a dict
    ``{name_in_checkpoint: variable}`` -- e.g. to load the ps copies ``g/a`` of a DOWNPOUR run into
    a single-process model whose variable is called ``a``."""

    def __init__(self, var_list=None, max_to_keep: int = 5,
                 extra_state_fn=None, restore_state_fn=None, write_format: str = "dtg"):
        if write_format not in ("dtg", "tf_bundle"):
            raise ValueError("write_format must be 'dtg' or 'tf_bundle'")
        #: "tf_bundle": TF Saver V2 layout (<prefix>.index + .data-00000-of-00001, tf_bundle.py)
        self._write_format = write_format
        if isinstance(var_list, dict):
            self._named = [(str(k), v) for k, v in var_list.items()]
        else:
            self._named = None
        self._var_list = (list(var_list.values()) if isinstance(var_list, dict)
                          else list(var_list) if var_list is not None else None)
        self._max_to_keep = max_to_keep
        self._kept: List[str] = []
        #: engines with device-resident shards register a gather / scatter pair
        self._extra_state_fn = extra_state_fn
        self._restore_state_fn = restore_state_fn
        self.last_save_time = 0.0

    def _vars(self) -> List[Variable]:
        if self._var_list is not None:
            return self._var_list
        return fw.get_collection(GraphKeys.GLOBAL_VARIABLES)

    def _named_vars(self):
        """``[(name in the checkpoint, variable)]``."""
        if self._named is not None:
            return self._named
        return [(v.op.name, v) for v in self._vars()]

    @property
    def last_checkpoints(self) -> List[str]:
        """Prefixes of the checkpoints this saver still keeps, oldest first ([TF])."""
        return list(self._kept)

    def save(self, sess, save_path: str, global_step=None, latest_filename=None,
             write_state: bool = True) -> str:
        if global_step is not None:
            if isinstance(global_step, Variable):
                global_step = int(global_step.read_value().item())
            prefix = "%s-%d" % (save_path, int(global_step))
        else:
            prefix = save_path
        d = os.path.dirname(prefix) or "."
        os.makedirs(d, exist_ok=True)
        tensors = {}
        for name, v in self._named_vars():
            try:
                tensors[name] = v.read_value()
            except KeyError:
                continue  # uninitialised (e.g. slot not created yet)
        if self._extra_state_fn is not None:
            tensors.update(self._extra_state_fn())
        if self._write_format == "tf_bundle":
            from . import tf_bundle
            tf_bundle.write_bundle(prefix, tensors)
        else:
            save_tensors(prefix + SUFFIX, tensors)
        self.last_save_time = time.time()
        if prefix in self._kept:
            self._kept.remove(prefix)
        self._kept.append(prefix)
        while self._max_to_keep and len(self._kept) > self._max_to_keep:
            old = self._kept.pop(0)
            for ext in (SUFFIX, ".index", ".data-00000-of-00001"):
                try:
                    os.remove(old + ext)
                except OSError:
                    pass
        if write_state:
            update_checkpoint_state(d, prefix, self._kept, latest_filename)
        return prefix

    def restore(self, sess, save_path: str) -> List[str]:
        data = load_checkpoint(save_path)
        restored = []
        for name, v in self._named_vars():
            t = data.get(name)
            if t is not None:
                want = tuple(getattr(v, "shape", ()) or ())
                if want and tuple(t.shape) != want:
                    raise ValueError("checkpoint tensor %r has shape %s, variable %s wants %s" % (
                        name, tuple(t.shape), v.op.name, want))
                v.assign(t)
                restored.append(name)
        if self._restore_state_fn is not None:
            self._restore_state_fn(data)
        return restored

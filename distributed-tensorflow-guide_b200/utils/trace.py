"""Tracing for the scheme engines: NVTX ranges, device-timed phases, chrome-trace timelines and
roofline fractions ("Tracing / profiling" and "Metrics / logging").

The reference has none of this (only ``log_device_placement=False``,
/root/reference/DOWNPOUR/DOWNPOUR.py:21); TF-1 users reached for ``ProfilerHook``'s
``timeline-<step>.json``.  Here:

* ``PhaseTracer.range(name, nbytes=..., kind=...)`` brackets a phase of a step (pull, forward,
  backward+push, bucket sync, local steps ...) with an NVTX range -- visible to any NVTX-aware
  profiler -- and a pair of CUDA events recorded on the *current stream*, so side-stream
  phases (prefetched pull, bucketed reduce-scatter) are timed where they run.  Events are
  resolved lazily (``query()``, never a synchronise inside a step); nothing is recorded and no
  event is created while the tracer is disabled.
* ``summary()`` gives calls / mean / total ms per phase and, for phases that declared bytes or
  FLOPs, GB/s or TFLOP/s and the fraction of the matching roofline: NVLink bytes against the
  900 GB/s port, HBM bytes and bf16 FLOPs against ``MEASURED_PEAKS.json`` when present.
* ``write_chrome_trace(path)`` writes the timeline (one lane per stream) in the Trace Event
  format chrome://tracing and Perfetto read; ``write_jsonl(path)`` appends the summary as one
  line per phase.
* ``ProfilerHook(save_steps / save_secs, output_dir)`` does for a ``GpuTrainingSession`` (or any
  hooked session whose tracer is passed in) what ``tf.train.ProfilerHook`` did.

Without CUDA the same API times with ``time.perf_counter`` (host engines, CPU tests).
Enable from the environment with ``DTG_TRACE=1`` (``DTG_TRACE_DIR`` for the output directory).
"""
from __future__ import annotations

import json
import os
import time
from typing import Dict, List, Optional

import torch

__all__ = ["PhaseTracer", "ProfilerHook", "load_peaks", "NVLINK_PORT_GBPS"]

NVLINK_PORT_GBPS = 900.0           # per direction, per GPU (NVLink 5, 18 links)
_FALLBACK_PEAKS = {"hbm_gbs": 7700.0, "bf16_tflops": 2250.0, "bf16_tflops_sustained": 2250.0}


def load_peaks(root: Optional[str] = None) -> Dict[str, float]:
    """Roofline denominators: the driver-written ``MEASURED_PEAKS.json`` at the repo root when it
    exists, else the nominal B200 figures."""
    here = root or os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    out = dict(_FALLBACK_PEAKS)
    out["source"] = "nominal"
    try:
        with open(os.path.join(here, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        for k in ("hbm_gbs", "bf16_tflops", "bf16_tflops_sustained"):
            if isinstance(m.get(k), (int, float)) and m[k] > 0:
                out[k] = float(m[k])
        out["source"] = "MEASURED_PEAKS.json"
    except (OSError, ValueError):
        pass
    return out


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


_NULL = _Null()


class _Range:
    __slots__ = ("tr", "name", "nbytes", "kind", "flops", "lane", "t0", "e0")

    def __init__(self, tr, name, nbytes, kind, flops, lane):
        self.tr, self.name, self.nbytes, self.kind = tr, name, nbytes, kind
        self.flops, self.lane = flops, lane

    def __enter__(self):
        tr = self.tr
        if tr.nvtx:
            torch.cuda.nvtx.range_push(self.name)
        if tr.cuda:
            self.e0 = tr._new_event()
            if tr._base is None:
                tr._base = tr._new_event()
                tr._base.record(torch.cuda.current_stream())
            self.e0.record(torch.cuda.current_stream())
        else:
            self.t0 = tr._clock()
        return self

    def __exit__(self, *exc):
        tr = self.tr
        if tr.cuda:
            e1 = tr._new_event()
            e1.record(torch.cuda.current_stream())
            tr._pending.append((self.name, tr.step, self.e0, e1, self.nbytes, self.kind,
                                self.flops, self.lane))
        else:
            t1 = tr._clock()
            if tr._t_base is None:
                tr._t_base = self.t0
            tr._add(self.name, tr.step, (self.t0 - tr._t_base) * 1e3, (t1 - self.t0) * 1e3,
                    self.nbytes, self.kind, self.flops, self.lane)
        if tr.nvtx:
            torch.cuda.nvtx.range_pop()
        return False


class PhaseTracer:
    """See the module docstring.  ``kind``: ``"link"`` (bytes crossed NVLink), ``"hbm"`` (bytes
    moved through this GPU's HBM) or ``"compute"`` (``flops`` given)."""

    def __init__(self, enabled: bool = True, device=None, rank: int = 0,
                 nvtx: Optional[bool] = None, max_records: int = 200000, clock=None,
                 peaks: Optional[Dict[str, float]] = None):
        self.enabled = bool(enabled)
        dev = torch.device(device) if device is not None else None
        self.cuda = bool(dev is not None and dev.type == "cuda" and torch.cuda.is_available())
        self.nvtx = self.cuda if nvtx is None else bool(nvtx and self.cuda)
        self.rank = int(rank)
        self.step = 0
        self.max_records = max_records
        self.peaks = peaks or load_peaks()
        self._clock = clock or time.perf_counter
        self._t_base: Optional[float] = None
        self._base = None
        self._free: List = []
        self._pending: List = []
        self.records: List[dict] = []
        self.dropped = 0

    # ------------------------------------------------------------------ setup
    @classmethod
    def from_env(cls, device=None, rank: int = 0) -> "PhaseTracer":
        on = os.environ.get("DTG_TRACE", "0") not in ("", "0")
        return cls(enabled=on, device=device, rank=rank)

    def _new_event(self):
        return self._free.pop() if self._free else torch.cuda.Event(enable_timing=True)

    # ----------------------------------------------------------------- record
    def range(self, name: str, nbytes: int = 0, kind: str = "compute", flops: float = 0.0,
              lane: str = "main"):
        if not self.enabled:
            return _NULL
        return _Range(self, name, int(nbytes), kind, float(flops), lane)

    def _add(self, name, step, t0_ms, dur_ms, nbytes, kind, flops, lane):
        if len(self.records) >= self.max_records:
            self.dropped += 1
            return
        self.records.append({"name": name, "step": step, "t0_ms": t0_ms, "dur_ms": dur_ms,
                             "bytes": nbytes, "kind": kind, "flops": flops, "lane": lane})

    def step_done(self) -> None:
        """End of one engine step: number the next one and resolve what has already finished
        (``query()`` only -- no synchronise on the training path)."""
        if not self.enabled:
            return
        self.step += 1
        self._resolve(block=False)

    def _resolve(self, block: bool) -> None:
        if not self.cuda or not self._pending:
            return
        if block:
            torch.cuda.synchronize()
        keep = []
        for rec in self._pending:
            name, step, e0, e1, nbytes, kind, flops, lane = rec
            if not block and not e1.query():
                keep.append(rec)
                continue
            self._add(name, step, self._base.elapsed_time(e0), e0.elapsed_time(e1), nbytes, kind,
                      flops, lane)
            self._free.extend((e0, e1))
        self._pending = keep

    def collect(self) -> List[dict]:
        """Synchronise and return every record so far."""
        self._resolve(block=True)
        return self.records

    def reset(self) -> None:
        self.collect()
        self.records, self.dropped = [], 0

    # ---------------------------------------------------------------- reports
    def summary(self, skip_steps: int = 0) -> Dict[str, dict]:
        """Per phase: calls, total / mean / max ms, and for phases with declared traffic or work
        the achieved rate and its fraction of the roofline.  ``skip_steps`` drops warm-up."""
        out: Dict[str, dict] = {}
        for r in self.collect():
            if r["step"] < skip_steps:
                continue
            s = out.setdefault(r["name"], {"calls": 0, "ms_total": 0.0, "ms_max": 0.0, "bytes": 0,
                                           "flops": 0.0, "kind": r["kind"], "lane": r["lane"]})
            s["calls"] += 1
            s["ms_total"] += r["dur_ms"]
            s["ms_max"] = max(s["ms_max"], r["dur_ms"])
            s["bytes"] += r["bytes"]
            s["flops"] += r["flops"]
        for s in out.values():
            s["ms_mean"] = s["ms_total"] / max(1, s["calls"])
            ms = s["ms_total"]
            if s["bytes"] and ms > 0:
                s["GBps"] = s["bytes"] / ms / 1e6
                peak = NVLINK_PORT_GBPS if s["kind"] == "link" else self.peaks["hbm_gbs"]
                s["roofline"] = "nvlink_port_900" if s["kind"] == "link" else "hbm_" + self.peaks["source"]
                s["fraction_of_roofline"] = s["GBps"] / peak
            if s["flops"] and ms > 0:
                s["TFLOPs"] = s["flops"] / ms / 1e9
                s["fraction_of_bf16_peak"] = s["TFLOPs"] / self.peaks["bf16_tflops_sustained"]
        return out

    def write_chrome_trace(self, path: str, skip_steps: int = 0) -> str:
        """Trace Event format: one process per rank, one thread (lane) per stream."""
        lanes: Dict[str, int] = {}
        ev = []
        for r in self.collect():
            if r["step"] < skip_steps:
                continue
            tid = lanes.setdefault(r["lane"], len(lanes))
            args = {"step": r["step"]}
            if r["bytes"]:
                args["bytes"] = r["bytes"]
                if r["dur_ms"] > 0:
                    args["GBps"] = round(r["bytes"] / r["dur_ms"] / 1e6, 1)
            ev.append({"name": r["name"], "ph": "X", "pid": self.rank, "tid": tid,
                       "ts": r["t0_ms"] * 1e3, "dur": r["dur_ms"] * 1e3, "cat": r["kind"],
                       "args": args})
        meta = [{"name": "process_name", "ph": "M", "pid": self.rank,
                 "args": {"name": "rank %d" % self.rank}}]
        meta += [{"name": "thread_name", "ph": "M", "pid": self.rank, "tid": t,
                  "args": {"name": "stream:" + n}} for n, t in lanes.items()]
        d = os.path.dirname(os.path.abspath(path))
        os.makedirs(d, exist_ok=True)
        with open(path, "w") as f:
            json.dump({"traceEvents": meta + ev, "displayTimeUnit": "ms"}, f)
        return path

    def write_jsonl(self, path: str, skip_steps: int = 0, **extra) -> str:
        d = os.path.dirname(os.path.abspath(path))
        os.makedirs(d, exist_ok=True)
        with open(path, "a") as f:
            for name, s in sorted(self.summary(skip_steps).items()):
                row = {"phase": name, "rank": self.rank}
                row.update({k: (round(v, 6) if isinstance(v, float) else v) for k, v in s.items()})
                row.update(extra)
                f.write(json.dumps(row) + "\n")
        return path


def _hook_base():
    from ..session.hooks import SessionRunHook
    return SessionRunHook


class ProfilerHook(_hook_base()):
    """``tf.train.ProfilerHook`` for the engines: every ``save_steps`` runs (or ``save_secs``
    seconds) write ``output_dir/timeline-<step>.json`` with the phases traced since the last
    dump and append their summary to ``output_dir/phases.jsonl``.  The tracer is switched on in
    ``begin()``; pass the engine's (``GpuWorker.tracer``)."""

    def __init__(self, tracer: PhaseTracer, save_steps: Optional[int] = None,
                 save_secs: Optional[float] = None, output_dir: str = "."):
        if (save_steps is None) == (save_secs is None):
            raise ValueError("exactly one of save_steps and save_secs should be provided.")
        from ..session.hooks import SecondOrStepTimer
        self._tracer, self._dir = tracer, output_dir
        self._timer = SecondOrStepTimer(every_secs=save_secs, every_steps=save_steps)
        self._runs = 0
        self.written: List[str] = []

    def begin(self):
        self._tracer.enabled = True

    def after_run(self, run_context, run_values):
        self._runs += 1
        if self._timer.should_trigger_for_step(self._runs):
            self._timer.update_last_triggered_step(self._runs)
            self._dump()

    def _dump(self):
        tr = self._tracer
        if not tr.collect():
            return
        p = os.path.join(self._dir, "timeline-%d.json" % self._runs)
        self.written.append(tr.write_chrome_trace(p))
        tr.write_jsonl(os.path.join(self._dir, "phases.jsonl"), run=self._runs)
        tr.records = []

    def end(self, session):
        self._dump()

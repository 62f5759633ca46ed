"""Scheme engines on B200: the same algorithms as ``engines.py``
with the data plane entirely in sm_100a kernels over NVLink peer memory.

A *native model* exposes flat buffers (``p32``, ``p16``, ``g32``, ``layout``) and
``forward_backward_native(x, y, sink)`` (``models/mlp_native.py`` is the
template).  One process per GPU; ``Fabric`` maps every rank's buffers.

============  ==============================================================
hogwild        pull (peer loads, fp32->bf16) -> fwd -> bwd whose wgrad epilogues
multigpu_async ``red.add`` ``-lr*dW`` into the ps shards -> ``global_step`` bump.  No
               barrier, no lock (Hogwild/Hogwild.py:1-8).
downpour       T local steps (fused Adagrad/SGD + window accumulate kernel), then
downpour_easy  one push of the window (Adagrad-on-ps: atom on acc, red on w; ADAG:
adag           ``-lr/T`` scaled red), then pull (DOWNPOUR/DOWNPOUR.py:63-102).
ssgd, ssgd_lr  gradients land (scaled by the worker's lr ratio) in a symmetric
tower_avg      buffer; one fused reduce-scatter + SGD/Adagrad + all-gather kernel
sdag           per bucket, launched on a side stream as soon as that bucket's
               wgrads are issued, overlapping the rest of backward.
============  ==============================================================
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch

from .. import ops
from .engines import SchemeConfig
from .fabric import Fabric
from .peer_store import PeerStore, PeerSync
from ..utils.trace import PhaseTracer

__all__ = ["GpuWorker"]

ASYNC = ("hogwild", "multigpu_async", "single")
WINDOW = ("downpour", "downpour_easy", "adag")
SYNC = ("ssgd", "ssgd_lr", "tower_avg", "sdag")
ELASTIC = ("aeasgd", "aeamsgd")           # + "dynsgd": the reference's TODO list (README.md:40-42)


class GpuWorker:
    def __init__(self, model, cfg: SchemeConfig, fabric: Fabric, task_index: Optional[int] = None,
                 n_workers: Optional[int] = None, ps_ranks: Optional[Sequence[int]] = None,
                 fused_push: bool = True, overlap: bool = True, seed: int = 0,
                 prefetch_pull: bool = True, striped_ps: bool = False,
                 async_push: Optional[bool] = None, tracer: Optional[PhaseTracer] = None,
                 device_init: bool = False):
        self.model, self.fabric = model, fabric
        # NVTX ranges + device-timed phases (utils/trace.py); off unless DTG_TRACE=1 or a
        # tracer / ProfilerHook switches it on -- a disabled tracer records nothing
        self.tracer = tracer if tracer is not None else PhaseTracer.from_env(fabric.device, fabric.rank)
        self.task_index = fabric.rank if task_index is None else task_index
        self.n_workers = fabric.world if n_workers is None else n_workers
        self.cfg = cfg.resolved(self.n_workers)
        self.is_chief = self.task_index == 0
        self.fused_push, self.overlap = fused_push, overlap
        self.rounds = 0
        s = self.cfg.scheme
        self.store: Optional[PeerStore] = None
        self.sync: Optional[PeerSync] = None
        if device_init and hasattr(model, "init_on_device"):
            model.init_on_device(seed)               # Philox on the GPU: same seed, same bits
        else:
            init = model.init_params(seed=seed)      # same seed on every rank
            model.load_params(init)
        if s in ASYNC or s in WINDOW or s in ELASTIC or s == "dynsgd":
            self.store = PeerStore(fabric, model.layout, None if striped_ps else ps_ranks,
                                   adagrad=s in ("downpour", "downpour_easy"), striped=striped_ps)
            if self.is_chief:
                self.store.seed(model.p32)            # chief seeds the ps (K6)
            self._push_sink = model.sink("push", self.store.sink_views(), alpha=-self.cfg.lr,
                                         remote=self.store.remote_names())
        else:
            bounds = None
            if overlap and hasattr(model, "layer_bounds"):
                lb = model.layer_bounds()
                cuts = sorted(set([0, model.layout.total] + [b for b, _ in lb]))
                bounds = cuts
                self._layer_bucket = {}
                for li, (lo, hi) in enumerate(lb):
                    self._layer_bucket[li] = max(i for i, c in enumerate(cuts[:-1]) if c <= lo)
            self.sync = PeerSync(fabric, model.layout.total, self.cfg.lr,
                                 replicas_to_aggregate=self.cfg.replicas_to_aggregate,
                                 bucket_bounds=bounds)
            model.rebind(self.sync.p32, self.sync.p16, self.sync.g32)
            self.sync.load_master(model.p32)
            scale = 1.0
            if s == "ssgd_lr":
                scale = (self.cfg.worker_lrs[self.task_index] / self.cfg.lr) * \
                    self.cfg.replicas_to_aggregate
            self._sync_sink = model.sink("store", model.g32, alpha=scale)
            self._comm_stream = torch.cuda.Stream(device=fabric.device)
        # a ring of reusable events (record / wait_event pairs are consumed in program order,
        # so re-recording an event whose wait has already been enqueued is safe)
        self._ev_pool: List[torch.cuda.Event] = [torch.cuda.Event() for _ in range(64)]
        self._ev_next = 0
        self._prefetch = (s in ASYNC and prefetch_pull and hasattr(model, "enable_double_buffer")
                          and hasattr(model, "set_input"))
        # Single-port ps (every shard on one GPU, BASELINE config 2 as written): a push fused into
        # the wgrad epilogues is throttled to that port's ingress rate while the tensor cores wait,
        # so the gradients are stored locally and pushed layer by layer on a side stream, where
        # the transfer overlaps ALL of the step's compute.  Striped / sharded ps: the fused push.
        single_port = (self.store is not None and not striped_ps and fabric.world > 2 and
                       len(set(self.store.ps_ranks)) == 1 and len(model.layout.shard_ranges) == 1)
        self.async_push = (single_port if async_push is None else bool(async_push)) and \
            s in ASYNC and hasattr(model, "layer_bounds") and self._prefetch_possible(model, prefetch_pull)
        if self.async_push:
            self._push_stream = torch.cuda.Stream(device=fabric.device)
            self._push_done: Optional[torch.cuda.Event] = None
            self._layer_ranges = model.layer_bounds()
        # Hogwild never steps locally: its pulls keep fp32 only where the model reads fp32
        self._pull_f32 = model.f32_ranges() if (s in ASYNC and hasattr(model, "f32_ranges")) else None
        if self._prefetch:
            model.enable_double_buffer()
            self._pull_stream = torch.cuda.Stream(device=fabric.device)
            self._pull_ready = None
        if s in ELASTIC:
            self._velocity = torch.zeros_like(model.p32)
        if s == "dynsgd":
            self._seen = torch.zeros(1, dtype=torch.int64, device=fabric.device)
            self.staleness = torch.ones(1, dtype=torch.int64, device=fabric.device)
        if s in WINDOW or s in ("sdag", "dynsgd"):
            self._window = torch.zeros_like(model.p32)
            self._slot = torch.full_like(model.p32, 0.1) if s == "downpour" else None
        # traffic / work declared to the tracer: bytes of one full pull or push that cross
        # NVLink (the part of the ps that is not in this GPU's HBM), FLOPs of one fwd+bwd
        pbytes = model.layout.total * 4
        if self.store is not None and striped_ps:
            self._link_bytes = pbytes * (fabric.world - 1) // fabric.world
        elif self.store is not None:
            self._link_bytes = 4 * sum(e - b for r, (b, e) in zip(self.store.ps_ranks,
                                                                  model.layout.shard_ranges)
                                       if r != fabric.rank)
        else:
            self._link_bytes = pbytes * 10 // 4 * (fabric.world - 1) // fabric.world
        B = getattr(model, "B", None)
        self._step_flops = float(model.flops_per_sample() * B) if (
            B and hasattr(model, "flops_per_sample")) else 0.0
        torch.cuda.synchronize(fabric.device)
        fabric.host_barrier()                         # every shard seeded before anyone pulls
        if self.store is not None:
            self.store.pull(model.p32, model.p16)     # "grab_global_init" (DOWNPOUR.py:135)

    @staticmethod
    def _prefetch_possible(model, prefetch_pull) -> bool:
        return bool(prefetch_pull) and hasattr(model, "enable_double_buffer") and hasattr(model, "set_input")

    def _push_layer_async(self, li: int) -> None:
        """Layer ``li``'s gradients are in ``g32``: add ``-lr * g`` into the ps on the push stream."""
        lo, hi = self._layer_ranges[li]
        ev = self._event()
        ev.record(torch.cuda.current_stream())
        self._push_stream.wait_event(ev)
        with torch.cuda.stream(self._push_stream):
            with self.tracer.range("push(layer)", (hi - lo) * 4 if self._link_bytes else 0, "link",
                                   lane="push"):
                ops.push_axpy(self.store.shards[0][lo:hi], self.model.g32[lo:hi], -self.cfg.lr)

    def _event(self) -> torch.cuda.Event:
        ev = self._ev_pool[self._ev_next]
        self._ev_next = (self._ev_next + 1) % len(self._ev_pool)
        return ev

    def _micro_batches(self, x, y):
        """The T (x, y) pairs of one communication window.

        ``x`` / ``y`` may be a sequence of T tensors, or one tensor holding T micro-batches
        stacked along dim 0 (``T * B`` rows: what ``bench.py`` and the input pipeline feed, so
        every local step sees fresh data as in the host engine's ``batch_fn()``).  A single
        batch of ``B`` rows is reused for all T local steps (the reference's toy problem has
        no input at all, ADAG/ADAG.py:69-83)."""
        T = self.cfg.update_window
        if isinstance(x, (list, tuple)):
            if len(x) != T or len(y) != T:
                raise ValueError("need %d micro-batches per window, got %d" % (T, len(x)))
            return list(zip(x, y))
        B = getattr(self.model, "B", None)            # the native models' per-step batch
        if B is not None and T > 1 and x.shape[0] == T * B:
            yb = y.shape[0] // T
            return [(x[t * B:(t + 1) * B], y[t * yb:(t + 1) * yb]) for t in range(T)]
        return [(x, y)] * T

    # -------------------------------------------------------------- schemes
    def step(self, x, y) -> torch.Tensor:
        """One ``sess.run(train_op)`` worth of work; returns the device loss tensor."""
        s = self.cfg.scheme
        self.rounds += 1
        if s in ASYNC:
            loss = self._step_hogwild(x, y)
        elif s in WINDOW:
            loss = self._step_window(x, y)
        elif s == "sdag":
            loss = self._step_sdag(x, y)
        elif s == "dynsgd":
            loss = self._step_dynsgd(x, y)
        elif s in ELASTIC:
            loss = self._step_elastic(x, y)
        else:
            loss = self._step_sync(x, y)
        self.tracer.step_done()
        return loss

    def _step_hogwild_prefetch(self, x, y):
        """Hogwild with the pull of step t+1 issued on a side stream as soon as step t's
        forward is done, so it overlaps backward.  The worker then computes on parameters
        that miss (at most) its own last push -- staleness Hogwild permits by definition."""
        m, st, tr = self.model, self.store, self.tracer
        main = torch.cuda.current_stream()
        if self._pull_ready is not None:
            main.wait_event(self._pull_ready)       # parameters for this step have landed
            m.swap_params()
        else:
            with tr.range("pull", self._link_bytes, "link"):
                st.pull(m.p32, m.p16, self._pull_f32)
        m.set_input(x)
        with tr.range("forward", flops=self._step_flops / 3):
            m.forward(y)
        fwd_done = self._event()
        fwd_done.record(main)
        self._pull_stream.wait_event(fwd_done)      # the alternate pair is free once fwd ran
        with torch.cuda.stream(self._pull_stream):
            with tr.range("pull(prefetch)", self._link_bytes, "link", lane="pull"):
                st.pull(m._alt[0], m._alt[1], self._pull_f32)
            self._pull_ready = self._event()
            self._pull_ready.record(self._pull_stream)
        if self.async_push:
            if self._push_done is not None:
                main.wait_event(self._push_done)    # last step's pushes have read g32
            m.after_layer = self._push_layer_async
            with tr.range("backward", flops=2 * self._step_flops / 3):
                m.backward(m.local_sink)
            m.after_layer = None
            self._push_done = self._event()
            self._push_done.record(self._push_stream)
        elif self.fused_push:
            with tr.range("backward+push", self._link_bytes, "link", 2 * self._step_flops / 3):
                m.backward(self._push_sink)
        else:
            with tr.range("backward", flops=2 * self._step_flops / 3):
                m.backward(m.local_sink)
            with tr.range("push", self._link_bytes, "link"):
                st.push_sgd(m.g32, -self.cfg.lr)
        st.bump_step()
        return m.loss_buf

    def _step_hogwild(self, x, y):
        if self._prefetch:
            return self._step_hogwild_prefetch(x, y)
        m, st, tr = self.model, self.store, self.tracer
        with tr.range("pull", self._link_bytes, "link"):
            st.pull(m.p32, m.p16, self._pull_f32)
        if self.fused_push:
            with tr.range("fwd+bwd+push", self._link_bytes, "link", self._step_flops):
                loss = m.forward_backward_native(x, y, self._push_sink)
        else:
            with tr.range("fwd+bwd", flops=self._step_flops):
                loss = m.forward_backward_native(x, y)
            with tr.range("push", self._link_bytes, "link"):
                st.push_sgd(m.g32, -self.cfg.lr)
        st.bump_step()
        return loss

    def _local_rule(self):
        s = self.cfg.scheme
        return s == "downpour", self.cfg.local_lr_for(self.task_index)

    def _step_window(self, x, y):
        m, st, c, tr = self.model, self.store, self.cfg, self.tracer
        adagrad, llr = self._local_rule()
        self._window.zero_()
        loss = None
        with tr.range("local_steps", flops=self._step_flops * c.update_window):
            for xb, yb in self._micro_batches(x, y):
                loss = m.forward_backward_native(xb, yb)
                ops.local_step(m.p32, m.g32, llr, p_bf16=m.p16, slot=self._slot,
                               window=self._window, adagrad=adagrad)
        if c.scheme == "adag":
            with tr.range("push", self._link_bytes, "link"):
                st.push_sgd(self._window, -c.lr / c.update_window)   # normalised (mean)
        else:
            with tr.range("push_adagrad", 2 * self._link_bytes, "link"):
                st.push_adagrad(self._window, c.lr)                   # sum, Adagrad on ps
        st.bump_step()
        with tr.range("pull", self._link_bytes, "link"):
            st.pull(m.p32, m.p16)
        return loss

    def _step_dynsgd(self, x, y):
        """Staleness-aware async SGD: pull (remember the ps version), T local SGD steps with a
        window sum, push weighted by 1/staleness -- all counters stay on the device."""
        m, st, c = self.model, self.store, self.cfg
        st.pull(m.p32, m.p16)
        st.snapshot_step(self._seen)
        self._window.zero_()
        loss = None
        llr = c.local_lr_for(self.task_index)
        for xb, yb in self._micro_batches(x, y):
            loss = m.forward_backward_native(xb, yb)
            ops.local_step(m.p32, m.g32, llr, p_bf16=m.p16, window=self._window)
        st.push_sgd_dyn(self._window, -c.lr, self._seen, self.staleness)
        st.bump_step()
        return loss

    def _step_elastic(self, x, y):
        """AEASGD / AEAMSGD: one fused pull+push elastic exchange with the centre on the ps,
        then tau local (Nesterov-)SGD steps on the worker's own parameters."""
        from ..ops import elastic
        m, st, c = self.model, self.store, self.cfg
        delta = c.momentum if c.scheme == "aeamsgd" else 0.0
        st.elastic_exchange(m.p32, m.p16, c.elastic_alpha, self._velocity if delta else None, delta)
        st.bump_step()
        llr = c.local_lr_for(self.task_index)
        loss = None
        for xb, yb in self._micro_batches(x, y):
            loss = m.forward_backward_native(xb, yb)
            elastic.momentum_step(m.p32, self._velocity, m.g32, m.p16, llr, delta)
        return loss

    def _launch_bucket(self, bi: int):
        main = torch.cuda.current_stream()
        ev = self._event()
        ev.record(main)
        self._comm_stream.wait_event(ev)
        with torch.cuda.stream(self._comm_stream):
            bk = self.sync.buckets[bi]
            frac = (bk["end"] - bk["base"]) / max(1, self.model.layout.total)
            with self.tracer.range("sync_bucket", int(self._link_bytes * frac), "link", lane="comm"):
                self.sync.step_bucket(bi)

    def _step_sync(self, x, y):
        m, sy = self.model, self.sync
        sy.begin_step()
        nb = len(sy.buckets)
        if self.overlap and nb > 1:
            issued = set()
            remaining = {}
            for li, bi in self._layer_bucket.items():
                remaining[bi] = remaining.get(bi, 0) + 1

            def after_layer(li):
                bi = self._layer_bucket[li]
                remaining[bi] -= 1
                if remaining[bi] == 0:
                    self._launch_bucket(bi)
                    issued.add(bi)
            m.after_layer = after_layer
            with self.tracer.range("fwd+bwd", flops=self._step_flops):
                loss = m.forward_backward_native(x, y, self._sync_sink)
            m.after_layer = None
            for bi in range(nb):
                if bi not in issued:
                    self._launch_bucket(bi)
            done = self._event()
            done.record(self._comm_stream)
            torch.cuda.current_stream().wait_event(done)
        else:
            with self.tracer.range("fwd+bwd", flops=self._step_flops):
                loss = m.forward_backward_native(x, y, self._sync_sink)
            with self.tracer.range("sync", self._link_bytes, "link"):
                for bi in range(nb):
                    sy.step_bucket(bi)
        return loss

    def _step_sdag(self, x, y):
        m, sy, c = self.model, self.sync, self.cfg
        llr = c.local_lr_for(self.task_index)
        self._window.zero_()
        loss = None
        for xb, yb in self._micro_batches(x, y):
            loss = m.forward_backward_native(xb, yb)
            ops.local_step(m.p32, m.g32, llr, p_bf16=m.p16, window=self._window)
        m.g32.copy_(self._window)
        sy.begin_step()
        for bi in range(len(sy.buckets)):
            sy.step_bucket(bi, gscale=1.0 / c.update_window)
        return loss

    # ---------------------------------------------------------------- queries
    def global_step(self) -> int:
        if self.store is not None:
            return self.store.read_step()
        return self.sync.step_no

    def status(self) -> int:
        if self.sync is not None:
            return int(self.sync.status.item())
        return 0

    def named_state(self, prefix: str = "") -> Dict[str, torch.Tensor]:
        """Checkpoint payload keyed by TF variable name (chief calls this)."""
        if self.store is not None:
            return self.store.gather_named(prefix)
        full = self.sync.gather_master().detach().cpu()
        out = {prefix + n: v.clone() for n, v in self.model.layout.unpack(full).items()}
        out["global_step"] = torch.tensor(self.sync.step_no, dtype=torch.int64)
        return out

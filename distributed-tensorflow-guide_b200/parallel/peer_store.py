"""B200 parameter store: ps shards in NVLink-mapped HBM, touched by worker kernels.

Same role as ``store.HostStore`` but the data plane is device code:

=====================  ======================================================
reference (TF, gRPC)    here
=====================  ======================================================
variable read (K5)      ``ops.pull_convert``: peer loads fp32 -> local fp32+bf16
ApplyGradientDescent    ``red.global.add.v4.f32`` into the shard -- either the
on ps (K3)              standalone ``ops.push_axpy`` or fused in the wgrad GEMM
                        epilogue (``sink('push')``)
ApplyAdagrad on ps (K4) ``ops.push_adagrad``: ``atom.add`` on the accumulator,
                        ``red.add`` on the weight
assign_local_to_global  chief seeds the shards with a peer copy (K6)
global_step AssignAdd   ``red.global.add.u64`` on a word in shard 0 (K14)
SyncReplicas (K9-K12)   ``PeerSync``: one fused reduce-scatter + update +
                        all-gather kernel with flag-based token release
=====================  ======================================================

Shard ``s`` of the flat layout (``parallel/flat.py``) lives on fabric rank
``ps_ranks[s]``.  With ``ps_ranks = [0]`` this is BASELINE.json config 2
("ps shard on GPU0"); with ``ps_ranks = range(world)`` the ps is sharded over all
GPUs (config 4) and no single NVLink port is the hot spot.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence

import torch

from .. import ops
from ..placement import flat_shards
from .fabric import Fabric
from .flat import FlatLayout

__all__ = ["PeerStore", "PeerSync", "SyncParams"]


class PeerStore:
    def __init__(self, fabric: Fabric, layout: FlatLayout, ps_ranks: Optional[Sequence[int]] = None,
                 adagrad: bool = False, initial_accumulator_value: float = 0.1,
                 striped: bool = False):
        self.fabric, self.layout = fabric, layout
        self.striped = striped
        if striped and layout.num_shards != 1:
            raise ValueError("a striped ps is one flat shard (layout.num_shards == 1)")
        self.ps_ranks = list(ps_ranks) if ps_ranks is not None else \
            [s % fabric.world for s in range(layout.num_shards)]
        if len(self.ps_ranks) != layout.num_shards:
            raise ValueError("need one ps rank per shard")
        self.adagrad = adagrad
        self.shards: List[torch.Tensor] = []      # fp32 view of shard s (local or peer)
        self.accs: List[Optional[torch.Tensor]] = []
        for s, (b, e) in enumerate(layout.shard_ranges):
            owner = self.ps_ranks[s]
            n = max(e - b, 4)
            if striped:
                # one virtual range, 2 MB stripes rotating over every rank's HBM
                self.shards.append(fabric.alloc_striped("ps_striped", n * 4).tensor(torch.float32, 0, n))
                if adagrad:
                    t = fabric.alloc_striped("ps_striped_acc", n * 4).tensor(torch.float32, 0, n)
                    if fabric.rank == 0:
                        t.fill_(initial_accumulator_value)
                    self.accs.append(t)
                else:
                    self.accs.append(None)
                continue
            buf = fabric.alloc("ps_shard_%d" % s, n * 4 if fabric.rank == owner else 256)
            self.shards.append(buf.tensor(owner, torch.float32, 0, n))
            if adagrad:
                abuf = fabric.alloc("ps_acc_%d" % s, n * 4 if fabric.rank == owner else 256)
                t = abuf.tensor(owner, torch.float32, 0, n)
                if fabric.rank == owner:
                    t.fill_(initial_accumulator_value)
                self.accs.append(t)
            else:
                self.accs.append(None)
        sbuf = fabric.alloc("ps_step", 256)
        self.step_word = sbuf.tensor(self.ps_ranks[0], torch.int64, 0, 1)
        self._step_host = torch.zeros(1, dtype=torch.int64).pin_memory()
        self.bytes_pushed = 0
        self.bytes_pulled = 0

    # ------------------------------------------------------------------ init
    def seed(self, p32_local: torch.Tensor) -> None:
        """K6: chief's values -> every shard (peer memcpy)."""
        lib = ops.lib()
        for s, (b, e) in enumerate(self.layout.shard_ranges):
            if e > b and self.striped:
                ops.pull_convert(p32_local[b:e], None, self.shards[s][:e - b])   # stores, by kernel
            elif e > b:
                lib.dtg_memcpy_async(self.shards[s].data_ptr(), p32_local[b:e].data_ptr(),
                                     (e - b) * 4, torch.cuda.current_stream().cuda_stream)

    # ------------------------------------------------------------- data plane
    def pull(self, p32_local: torch.Tensor, p16_local: Optional[torch.Tensor],
             f32_ranges: Optional[Sequence] = None) -> None:
        """ps -> worker.  ``f32_ranges`` (flat element ranges): the only places the worker reads
        in fp32 (biases, norm parameters); everything else is kept in bf16 only, which drops the
        4 B/parameter fp32 write of the pull for schemes that never step locally (Hogwild)."""
        for s, (b, e) in enumerate(self.layout.shard_ranges):
            if e <= b:
                continue
            if f32_ranges is None or p16_local is None:
                ops.pull_convert(self.shards[s][:e - b], None if p16_local is None else p16_local[b:e],
                                 p32_local[b:e])
            else:
                ops.pull_convert(self.shards[s][:e - b], p16_local[b:e], None)
                part = [(max(lo, b) - b, min(hi, e) - b) for lo, hi in f32_ranges
                        if min(hi, e) > max(lo, b)]
                if part:
                    ops.pull_ranges(self.shards[s][:e - b], p32_local[b:e], part)
            self.bytes_pulled += (e - b) * 4

    def push_sgd(self, g32_local: torch.Tensor, alpha: float) -> None:
        for s, (b, e) in enumerate(self.layout.shard_ranges):
            if e > b:
                ops.push_axpy(self.shards[s][:e - b], g32_local[b:e], alpha)
                self.bytes_pushed += (e - b) * 4

    def push_adagrad(self, g32_local: torch.Tensor, lr: float, gscale: float = 1.0) -> None:
        if not self.adagrad:
            raise RuntimeError("store was created without Adagrad accumulators")
        for s, (b, e) in enumerate(self.layout.shard_ranges):
            if e > b:
                ops.push_adagrad(self.shards[s][:e - b], self.accs[s][:e - b], g32_local[b:e], lr,
                                 gscale)
                self.bytes_pushed += (e - b) * 8

    def push_sgd_dyn(self, g32_local: torch.Tensor, alpha: float, seen: torch.Tensor,
                     staleness_out: Optional[torch.Tensor] = None) -> None:
        """DynSGD: the push is weighted by 1/staleness, read from the ps step word on device."""
        from ..ops import elastic
        for s, (b, e) in enumerate(self.layout.shard_ranges):
            if e > b:
                elastic.push_axpy_dyn(self.shards[s][:e - b], g32_local[b:e], alpha, self.step_word,
                                      seen, staleness_out)
                self.bytes_pushed += (e - b) * 4

    def elastic_exchange(self, x32_local: torch.Tensor, x16_local: Optional[torch.Tensor],
                         alpha: float, velocity: Optional[torch.Tensor] = None,
                         delta: float = 0.0) -> None:
        """EASGD: pull and push fused -- ``e = alpha (x - centre); x -= e; centre += e``."""
        from ..ops import elastic
        for s, (b, e) in enumerate(self.layout.shard_ranges):
            if e > b:
                elastic.easgd_elastic(self.shards[s][:e - b], x32_local[b:e],
                                      None if x16_local is None else x16_local[b:e], alpha,
                                      None if velocity is None else velocity[b:e], delta)
                self.bytes_pulled += (e - b) * 4
                self.bytes_pushed += (e - b) * 4

    def snapshot_step(self, out: torch.Tensor) -> None:
        """Device copy of the ps step word (the version a pull saw)."""
        ops.lib().dtg_memcpy_async(out.data_ptr(), self.step_word.data_ptr(), 8,
                                   torch.cuda.current_stream().cuda_stream)

    def sink_views(self) -> Dict[str, torch.Tensor]:
        """Per-variable views into the (peer) shards: targets of the fused wgrad push."""
        out = {}
        for name in self.layout.order:
            e = self.layout.entries[name]
            off = self.layout.offset_in_shard(name)
            out[name] = self.shards[e.shard][off:off + e.numel].view(e.stored_shape)
        return out

    def remote_names(self):
        """Variables whose shard is owned by another rank (reached over NVLink)."""
        if self.striped:
            return set(self.layout.order) if self.fabric.world > 1 else set()
        return {n for n in self.layout.order
                if self.ps_ranks[self.layout.entries[n].shard] != self.fabric.rank}

    def bump_step(self, n: int = 1) -> None:
        ops.counter_add(self.step_word, n)

    def read_step(self) -> int:
        lib = ops.lib()
        lib.dtg_memcpy_async(self._step_host.data_ptr(), self.step_word.data_ptr(), 8,
                             torch.cuda.current_stream().cuda_stream)
        torch.cuda.current_stream().synchronize()
        return int(self._step_host.item())

    # ------------------------------------------------------------ checkpoints
    def gather_named(self, prefix: str = "") -> Dict[str, torch.Tensor]:
        """K15: D2H copy of every shard, split by TF variable name."""
        lib = ops.lib()
        out = {}
        for s, (b, e) in enumerate(self.layout.shard_ranges):
            if e <= b:
                continue
            host = torch.empty(e - b, dtype=torch.float32).pin_memory()
            src = self.shards[s]
            if self.striped:                      # stage through local HBM with a kernel
                src = torch.empty(e - b, device=self.fabric.device)
                ops.pull_convert(self.shards[s][:e - b], None, src)
            lib.dtg_memcpy_async(host.data_ptr(), src.data_ptr(), (e - b) * 4,
                                 torch.cuda.current_stream().cuda_stream)
            acc_host = None
            if self.adagrad:
                acc_host = torch.empty(e - b, dtype=torch.float32).pin_memory()
                asrc = self.accs[s]
                if self.striped:
                    asrc = torch.empty(e - b, device=self.fabric.device)
                    ops.pull_convert(self.accs[s][:e - b], None, asrc)
                lib.dtg_memcpy_async(acc_host.data_ptr(), asrc.data_ptr(), (e - b) * 4,
                                     torch.cuda.current_stream().cuda_stream)
            torch.cuda.current_stream().synchronize()
            for name in self.layout.order:
                ent = self.layout.entries[name]
                if ent.shard != s:
                    continue
                off = ent.offset - b
                sl = tuple(slice(0, d) for d in ent.shape)
                out[prefix + name] = host[off:off + ent.numel].view(ent.stored_shape)[sl].clone()
                if acc_host is not None:
                    out[prefix + name + "/Adagrad"] = \
                        acc_host[off:off + ent.numel].view(ent.stored_shape)[sl].clone()
        out["global_step"] = torch.tensor(self.read_step(), dtype=torch.int64)
        return out

    def restore_named(self, data: Dict[str, torch.Tensor], prefix: str = "") -> None:
        lib = ops.lib()
        for s, (b, e) in enumerate(self.layout.shard_ranges):
            if e <= b or self.fabric.rank != self.ps_ranks[s]:
                continue
            host = torch.zeros(e - b, dtype=torch.float32)
            acc_host = torch.full((e - b,), 0.1, dtype=torch.float32)
            for name in self.layout.order:
                ent = self.layout.entries[name]
                if ent.shard != s:
                    continue
                off = ent.offset - b
                sl = tuple(slice(0, d) for d in ent.shape)
                if prefix + name in data:
                    host[off:off + ent.numel].view(ent.stored_shape)[sl] = data[prefix + name]
                if prefix + name + "/Adagrad" in data:
                    acc_host[off:off + ent.numel].view(ent.stored_shape)[sl] = \
                        data[prefix + name + "/Adagrad"]
            if self.striped:
                ops.pull_convert(host.to(self.fabric.device), None, self.shards[s][:e - b])
                if self.adagrad:
                    ops.pull_convert(acc_host.to(self.fabric.device), None, self.accs[s][:e - b])
                continue
            self.shards[s][:e - b].copy_(host.to(self.fabric.device))
            if self.adagrad:
                self.accs[s][:e - b].copy_(acc_host.to(self.fabric.device))
        if "global_step" in data and self.fabric.rank == self.ps_ranks[0]:
            self.step_word.fill_(int(data["global_step"]))


# --------------------------------------------------------------------------- #
# synchronous path
# --------------------------------------------------------------------------- #
class SyncParams(ctypes.Structure):
    """Mirror of ``dtg::SyncParams`` (ops/csrc/sync_kernels.cu)."""
    _fields_ = [
        ("grad", ctypes.c_void_p * 8), ("param_f32", ctypes.c_void_p * 8),
        ("param_bf16", ctypes.c_void_p * 8), ("flags", ctypes.c_void_p * 8),
        ("mc_grad", ctypes.c_void_p), ("mc_param_f32", ctypes.c_void_p),
        ("mc_param_bf16", ctypes.c_void_p), ("master", ctypes.c_void_p),
        ("slot", ctypes.c_void_p), ("decision", ctypes.c_void_p),
        ("cta_counter", ctypes.c_void_p), ("status", ctypes.c_void_p),
        ("shard_begin", ctypes.c_longlong), ("shard_end", ctypes.c_longlong),
        ("timeout_cycles", ctypes.c_longlong),
        ("world", ctypes.c_int), ("rank", ctypes.c_int),
        ("ready_base", ctypes.c_int), ("done_base", ctypes.c_int),
        ("replicas_to_aggregate", ctypes.c_int), ("step", ctypes.c_uint),
        ("lr", ctypes.c_float), ("gscale", ctypes.c_float),
        ("adagrad", ctypes.c_int), ("contributes", ctypes.c_int),
        ("owner_mask", ctypes.c_uint),
    ]


class PeerSync:
    """SSGD / per-worker-LR / SDAG / tower averaging on device.

    Every rank owns an equal slice of the flat parameter buffer (fp32 master +
    optional Adagrad slot, i.e. the ps is sharded over all GPUs), writes its
    gradients into a symmetric fp32 buffer, and calls ``step()``: one kernel that
    reduces its slice over the ranks that arrived, updates the master copy and
    broadcasts the new bf16/fp32 parameters into every rank's buffers.

    ``buckets`` splits the flat buffer so that reduction of late layers can start
    while earlier layers are still in backward (the caller launches ``step_bucket``
    on a side stream as soon as a bucket's gradients are complete).
    """

    MAX_BUCKETS = 48

    def __init__(self, fabric: Fabric, total_elems: int, lr: float, adagrad: bool = False,
                 replicas_to_aggregate: Optional[int] = None,
                 bucket_bounds: Optional[Sequence[int]] = None, keep_f32_params: bool = True,
                 num_blocks: int = 1184, contributes: bool = True, first_step: int = 0):
        """``contributes=False`` makes this rank a dedicated ps task: it owns and updates a shard
        but pushes no gradient (the reference's ps process; with ``replicas_to_aggregate`` =
        number of workers this is SyncReplicas with the ps outside the replica set).
        ``first_step``: steps are numbered from ``first_step + 1``; a second ``PeerSync`` on the
        same fabric reuses the signal-pad slots and must continue the first one's numbering."""
        self.fabric, self.total, self.lr, self.adagrad = fabric, int(total_elems), lr, adagrad
        if ops.lib().dtg_sync_params_size() != ctypes.sizeof(SyncParams):
            raise ops.NativeLibraryError("SyncParams layout mismatch between Python and the library")
        W, r = fabric.world, fabric.rank
        self.replicas_to_aggregate = replicas_to_aggregate or W
        self.num_blocks = num_blocks
        n = self.total
        # NVLS: gradients are summed and parameters broadcast inside the NVSwitch when the
        # buffers can be bound to multicast objects (full participation only; the N-of-M
        # gate falls back to peer loads for a partial step)
        self.nvls = False
        if fabric.multicast_supported():
            try:
                self.gbuf = fabric.alloc_multicast("sync_grad", n * 4)
                self.p16buf = fabric.alloc_multicast("sync_p16", n * 2)
                self.p32buf = fabric.alloc_multicast("sync_p32", n * 4) if keep_f32_params else None
                self.nvls = True
            except ops.NativeLibraryError as e:        # driver refused: plain peer memory
                import warnings
                warnings.warn("NVLS multicast unavailable (%s); using peer loads/stores" % e)
        if not self.nvls:
            self.gbuf = fabric.alloc("sync_grad", n * 4)
            self.p16buf = fabric.alloc("sync_p16", n * 2)
            self.p32buf = fabric.alloc("sync_p32", n * 4) if keep_f32_params else None
        self.g32 = self.gbuf.local(torch.float32, 0, n)
        self.p16 = self.p16buf.local(torch.bfloat16, 0, n)
        self.p32 = self.p32buf.local(torch.float32, 0, n) if keep_f32_params else None
        bounds = list(bucket_bounds) if bucket_bounds else [0, n]
        if bounds[0] != 0 or bounds[-1] != n or len(bounds) - 1 > self.MAX_BUCKETS:
            raise ValueError("bad bucket bounds")
        self.buckets = []
        dev = fabric.device
        for bi in range(len(bounds) - 1):
            b0, b1 = bounds[bi], bounds[bi + 1]
            if b0 % 4:
                raise ValueError("bucket bounds must be multiples of 4 elements")
            sh = flat_shards(b1 - b0, W, align_elems=4)
            sb, se = sh[r]
            self.buckets.append({
                "base": b0, "end": b1, "shard": (sb, se),
                "owner_mask": sum(1 << q for q in range(W) if sh[q][1] > sh[q][0]),
                "master": torch.zeros(max(se - sb, 4), device=dev, dtype=torch.float32),
                "slot": torch.full((max(se - sb, 4),), 0.1, device=dev, dtype=torch.float32)
                if adagrad else None,
            })
        self._decision = torch.zeros(len(self.buckets), dtype=torch.int64, device=dev)
        self._counter = torch.zeros(len(self.buckets), dtype=torch.int32, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        self.step_no = int(first_step)
        self.contributes = bool(contributes)
        self._params: List[SyncParams] = []
        for bi, bk in enumerate(self.buckets):
            p = SyncParams()
            for q in range(8):
                if q < W:
                    p.grad[q] = self.gbuf.ptrs[q] + bk["base"] * 4
                    p.param_bf16[q] = self.p16buf.ptrs[q] + bk["base"] * 2
                    p.param_f32[q] = (self.p32buf.ptrs[q] + bk["base"] * 4) if self.p32buf else None
                    p.flags[q] = fabric.flags.ptrs[q]
                else:
                    p.grad[q] = p.param_bf16[q] = p.param_f32[q] = p.flags[q] = None
            p.mc_grad = p.mc_param_f32 = p.mc_param_bf16 = None
            if self.nvls:
                p.mc_grad = self.gbuf.mc_ptr + bk["base"] * 4
                p.mc_param_bf16 = self.p16buf.mc_ptr + bk["base"] * 2
                if self.p32buf is not None:
                    p.mc_param_f32 = self.p32buf.mc_ptr + bk["base"] * 4
            p.master = bk["master"].data_ptr()
            p.slot = bk["slot"].data_ptr() if adagrad else None
            p.decision = self._decision[bi:].data_ptr()
            p.cta_counter = self._counter[bi:].data_ptr()
            p.status = self.status.data_ptr()
            p.shard_begin, p.shard_end = bk["shard"]
            p.timeout_cycles = fabric.timeout_cycles
            p.world, p.rank = W, r
            p.ready_base = 64 + bi * 16
            p.done_base = 64 + bi * 16 + 8
            p.replicas_to_aggregate = self.replicas_to_aggregate
            p.lr, p.gscale = lr, 1.0
            p.adagrad, p.contributes = int(adagrad), int(self.contributes)
            p.owner_mask = bk["owner_mask"]
            self._params.append(p)

    def load_master(self, p32_full: torch.Tensor) -> None:
        """Initialise master shards and the local parameter mirrors from a full copy."""
        for bk in self.buckets:
            sb, se = bk["shard"]
            if se > sb:
                bk["master"][:se - sb].copy_(p32_full[bk["base"] + sb: bk["base"] + se])
        if self.p32 is not None:
            self.p32.copy_(p32_full)
        ops.cast_f32_bf16(p32_full, self.p16)

    def begin_step(self) -> int:
        self.step_no += 1
        return self.step_no

    def step_bucket(self, bi: int, lr: Optional[float] = None, gscale: float = 1.0) -> None:
        p = self._params[bi]
        p.step = self.step_no
        p.lr = self.lr if lr is None else lr
        p.gscale = gscale
        rc = ops.lib().dtg_sync_rs_update_ag(ctypes.byref(p), self.num_blocks,
                                             torch.cuda.current_stream().cuda_stream)
        ops._check(rc, "dtg_sync_rs_update_ag")

    def step(self, lr: Optional[float] = None, gscale: float = 1.0) -> int:
        """All buckets on the current stream; returns the new global step (token)."""
        self.begin_step()
        for bi in range(len(self.buckets)):
            self.step_bucket(bi, lr, gscale)
        return self.step_no

    def gather_master(self) -> torch.Tensor:
        """Full fp32 parameters as currently mirrored locally (after a step)."""
        return self.p32 if self.p32 is not None else self.p16.float()

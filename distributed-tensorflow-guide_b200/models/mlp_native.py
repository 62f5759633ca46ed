"""sm_100a implementation of the MLP family (BASELINE.json configs 1-2): every
matmul is the tcgen05/TMEM/TMA GEMM of ``ops``, with bias+ReLU fused into the
forward epilogue, the ReLU mask fused into the dgrad epilogue, and the *gradient
sink* fused into the wgrad epilogue:

``store``  fp32 store into the local flat gradient buffer (SSGD / tower: the buffer
           is the symmetric one the reduce-scatter kernel reads; ``alpha`` carries
           the per-worker learning-rate scale, K13)
``acc``    ``+=`` into the window accumulator (DOWNPOUR / ADAG / SDAG, K8)
``push``   ``red.global.add.v4.f32`` of ``-lr * dW`` straight into the ps shard --
           Hogwild's unlocked ``ApplyGradientDescent`` (Hogwild/Hogwild.py:44) with
           no gradient tensor ever materialised (K2+K3)

Parameters live in flat buffers (``parallel/flat.py``): fp32 (master/local copy)
and a bf16 mirror that the GEMMs read.  The last layer's 10 outputs are padded to
16 rows so every row pitch is a multiple of 16 B (TMA requirement).
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import torch

from .. import ops
from ..parallel.flat import FlatLayout
from .mlp import MLP

__all__ = ["NativeMLP", "GradSink"]


def _pad8(n: int) -> int:
    return -(-n // 8) * 8


class GradSink:
    """Where wgrad / bias-grad epilogues write.  ``targets[name]`` is a 2-D (kernel)
    or 1-D (bias) fp32 view -- local buffer or peer-mapped ps shard."""

    def __init__(self, mode: str, targets: Dict[str, torch.Tensor], alpha: float = 1.0,
                 remote=()):
        assert mode in ("store", "acc", "push")
        self.mode, self.targets, self.alpha = mode, targets, alpha
        #: names whose target lives in a peer GPU's HBM: their push is staged so that the
        #: reductions leave as 128 B NVLink payloads instead of scattered 16 B ones
        self.remote = set(remote)

    def zero_reduction_targets(self, extra=()) -> None:
        """'store' sinks: every 1-D target (bias / norm parameter gradients, summed with red.add)
        and the named 2-D ones start from zero -- one launch for all of them."""
        key = tuple(extra)
        z = getattr(self, "_zero", None)
        if z is None or z[0] != key:
            ts = [t for n, t in self.targets.items() if t.dim() == 1 or n in key]
            z = (key, ops.ZeroMany(ts))
            self._zero = z
        z[1]()

    @property
    def epi(self) -> int:
        return {"store": ops.EPI_F32, "acc": ops.EPI_F32_ACC, "push": ops.EPI_F32_RED}[self.mode]

    def epi_for(self, name: str) -> int:
        if self.mode == "push" and name in self.remote:
            return ops.EPI_F32_RED_COAL
        return self.epi


class NativeMLP(MLP):
    name = "mlp_native"

    def __init__(self, sizes: Sequence[int], batch_size: int, device="cuda", num_shards: int = 1,
                 sharding: str = "round_robin"):
        super().__init__(sizes, compute_dtype=torch.bfloat16)
        self.device = torch.device(device)
        self.B = int(batch_size)
        L = self.num_layers
        self.stored_out = [_pad8(self.sizes[i + 1]) for i in range(L)]
        specs = []
        for i in range(L):
            k, b = self.layer_names(i)
            specs.append((k, (self.sizes[i + 1], self.sizes[i]), (self.stored_out[i], self.sizes[i])))
            specs.append((b, (self.sizes[i + 1],), (self.stored_out[i],)))
        if self.sizes[0] % 8:
            raise ValueError("input width must be a multiple of 8 (TMA row pitch)")
        self.layout = FlatLayout(specs, num_shards, sharding)
        dev, n = self.device, self.layout.total
        self.p32 = torch.zeros(n, device=dev, dtype=torch.float32)
        self.p16 = torch.zeros(n, device=dev, dtype=torch.bfloat16)
        self.g32 = torch.zeros(n, device=dev, dtype=torch.float32)
        self.after_layer = None        # callback(layer_index) once a layer's grads are issued
        self._window: Optional[torch.Tensor] = None
        self._slot: Optional[torch.Tensor] = None
        # activations
        self.acts = [torch.zeros(self.B, self.sizes[0], device=dev, dtype=torch.bfloat16)]
        for i in range(L - 1):
            self.acts.append(torch.zeros(self.B, self.stored_out[i], device=dev, dtype=torch.bfloat16))
        self.logits = torch.zeros(self.B, self.stored_out[-1], device=dev, dtype=torch.float32)
        wmax = max(self.stored_out)
        self._dz = [torch.zeros(self.B, wmax, device=dev, dtype=torch.bfloat16) for _ in range(2)]
        self.loss_buf = torch.zeros(1, device=dev, dtype=torch.float32)
        self._bind_views()

    def _bind_views(self) -> None:
        L = self.num_layers
        self.w16 = [self.layout.view(self.p16, self.layer_names(i)[0]) for i in range(L)]
        self.b32 = [self.layout.view(self.p32, self.layer_names(i)[1]) for i in range(L)]
        self.local_sink = GradSink("store", self._views(self.g32))

    def rebind(self, p32: torch.Tensor, p16: torch.Tensor, g32: torch.Tensor) -> None:
        """Use externally owned flat buffers (e.g. the fabric's symmetric ones)."""
        p32.copy_(self.p32); p16.copy_(self.p16); g32.zero_()
        self.p32, self.p16, self.g32 = p32, p16, g32
        self._bind_views()

    def enable_double_buffer(self) -> None:
        """Second fp32/bf16 parameter buffer pair: an asynchronous worker can pull the next
        step's parameters into it on a side stream while this step's backward still reads
        the current pair (``swap_params`` flips them)."""
        if getattr(self, "_alt", None) is None:
            self._alt = (torch.empty_like(self.p32), torch.empty_like(self.p16))

    def swap_params(self) -> None:
        (self.p32, self.p16), self._alt = self._alt, (self.p32, self.p16)
        L = self.num_layers
        self.w16 = [self.layout.view(self.p16, self.layer_names(i)[0]) for i in range(L)]
        self.b32 = [self.layout.view(self.p32, self.layer_names(i)[1]) for i in range(L)]

    def f32_ranges(self):
        """Flat ranges the forward/backward reads in fp32 (the biases): all a Hogwild pull has to
        refresh in fp32, the weight matrices are only ever read through the bf16 mirror."""
        out = []
        for i in range(self.num_layers):
            e = self.layout.entries[self.layer_names(i)[1]]
            out.append((e.offset, e.offset + -(-e.numel // 64) * 64))
        return out

    def layer_bounds(self):
        """Flat-buffer element ranges per layer (kernel+bias), for bucketed reduction."""
        out = []
        for i in range(self.num_layers):
            k, b = self.layer_names(i)
            ek, eb = self.layout.entries[k], self.layout.entries[b]
            lo = min(ek.offset, eb.offset)
            hi = max(ek.offset + -(-ek.numel // 64) * 64, eb.offset + -(-eb.numel // 64) * 64)
            out.append((lo, hi))
        return out

    # ---------------------------------------------------------------- params
    def _views(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {n: self.layout.view(flat, n) for n in self.layout.order}

    def sink(self, mode: str, flat_or_views, alpha: float = 1.0, remote=()) -> GradSink:
        views = flat_or_views if isinstance(flat_or_views, dict) else self._views(flat_or_views)
        return GradSink(mode, views, alpha, remote)

    def load_params(self, values: Dict[str, torch.Tensor]) -> None:
        """Logical (unpadded) tensors -> flat fp32 + bf16 mirror."""
        self.p32.zero_()
        self.layout.pack(self.p32, {k: v.to(self.device) for k, v in values.items()})
        self.refresh_bf16()

    def refresh_bf16(self) -> None:
        ops.cast_f32_bf16(self.p32, self.p16)

    def init_on_device(self, seed: int = 0) -> int:
        """Draw every parameter on the GPU, straight into the flat fp32 buffer (``ops/init.py``:
        counter-based, so equal seeds give bit-identical replicas on every rank)."""
        from ..ops import init as dinit
        n = dinit.device_init(self, self.layout, self.p32, seed)
        self.refresh_bf16()
        return n

    def export_params(self) -> Dict[str, torch.Tensor]:
        return {k: v.clone() for k, v in self.layout.unpack(self.p32).items()}

    def export_grads(self) -> Dict[str, torch.Tensor]:
        return {k: v.clone() for k, v in self.layout.unpack(self.g32).items()}

    # ------------------------------------------------------------------ data
    def synthetic_batch(self, batch_size=None, device=None, generator=None):
        B = batch_size or self.B
        x = torch.randn(B, self.sizes[0], generator=generator).to(self.device)
        y = torch.randint(0, self.sizes[-1], (B,), generator=generator).to(self.device)
        return x, y

    def set_input(self, x: torch.Tensor) -> None:
        """fp32 (or bf16) batch -> bf16 activation buffer 0."""
        if x.dtype == torch.bfloat16:
            self.acts[0].copy_(x)
        else:
            ops.cast_f32_bf16(x.contiguous(), self.acts[0])

    # ------------------------------------------------------- forward/backward
    def forward(self, labels: torch.Tensor, want_grad: bool = True) -> None:
        L = self.num_layers
        for i in range(L - 1):
            ops.gemm(self.acts[i], self.w16[i], self.acts[i + 1], bias=self.b32[i],
                     act=ops.ACT_RELU)
        ops.gemm(self.acts[L - 1], self.w16[L - 1], self.logits, bias=self.b32[L - 1],
                 mode=ops.EPI_F32)
        self.loss_buf.zero_()
        dl = self._dz[0][:, :self.stored_out[-1]] if want_grad else None
        ops.softmax_xent(self.logits, labels, dl, self.loss_buf, self.sizes[-1], 1.0 / self.B)

    def backward(self, sink: GradSink) -> None:
        L = self.num_layers
        cur = 0
        for i in range(L - 1, -1, -1):
            kname, bname = self.layer_names(i)
            dz = self._dz[cur][:, :self.stored_out[i]]
            # wgrad: dW[out, in] = dz^T . a[i]   (both operands MN-major), fused sink
            ops.gemm(dz, self.acts[i], sink.targets[kname], a_mn=True, b_mn=True,
                     alpha=sink.alpha, mode=sink.epi_for(kname))
            if sink.mode == "store":
                sink.targets[bname].zero_()
            ops.bias_grad(dz, sink.targets[bname], sink.alpha)
            if i > 0:
                # dgrad: da[B, in] = dz . W  (W is [out, in] = [K, N] -> MN-major B), ReLU mask fused
                nxt = self._dz[cur ^ 1][:, :self.stored_out[i - 1]]
                ops.gemm(dz, self.w16[i], nxt, b_mn=True, mask=self.acts[i],
                         mask_mode=ops.MASK_RELU)
                cur ^= 1
            if self.after_layer is not None:
                self.after_layer(i)

    def forward_backward_native(self, x: torch.Tensor, labels: torch.Tensor,
                                sink: Optional[GradSink] = None) -> torch.Tensor:
        """One fwd+bwd; returns the device-resident loss (no sync)."""
        self.set_input(x)
        self.forward(labels)
        self.backward(sink or self.local_sink)
        return self.loss_buf

    # Model protocol (dict in / dict out) -- used by tests and the generic engines
    def forward_backward(self, params, batch):
        self.load_params(params)
        x, y = batch
        loss = self.forward_backward_native(x, y).clone()
        return loss[0], self.export_grads()

    # ---------------------------------------------------------- local training
    def local_apply(self, lr: float, adagrad: bool = False, window: bool = False) -> None:
        """K7+K8: fused local optimizer step over the flat buffers (+ window sum)."""
        if adagrad and self._slot is None:
            self._slot = torch.full_like(self.p32, 0.1)
        if window and self._window is None:
            self._window = torch.zeros_like(self.p32)
        ops.local_step(self.p32, self.g32, lr, p_bf16=self.p16, slot=self._slot if adagrad else None,
                       window=self._window if window else None, adagrad=adagrad)

    def train_step_local(self, x, y, lr: float) -> float:
        loss = self.forward_backward_native(x, y)
        self.local_apply(lr)
        return float(loss.item())

"""Host -> device input staging and device -> host result reads that do not stall the step.

The reference feeds nothing (its graphs hold constants); a real job feeds a batch per step and
reads the loss back.  Done naively (copy, step, copy, synchronize) the PCIe transfers and the
host round trip sit on the critical path of every step.  ``DeviceFeeder`` stages batch ``i+1``
from pinned host memory on a copy stream while step ``i`` computes (double-buffered device
slots, slot reuse ordered by events); ``ResultReader`` copies each step's loss into a pinned
ring asynchronously and hands it to the host one step later, so the CPU keeps launching work.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

__all__ = ["DeviceFeeder", "ResultReader"]


class DeviceFeeder:
    def __init__(self, device, depth: int = 2):
        self.device = torch.device(device)
        self.depth = max(2, depth)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self._slots: List[Optional[Tuple[torch.Tensor, ...]]] = [None] * self.depth
        self._ready = [torch.cuda.Event() for _ in range(self.depth)]    # H2D of the slot finished
        self._free = [None] * self.depth                                  # consumer done with the slot
        self._put = 0
        self._got = 0
        self.bytes_per_put = 0

    def put(self, *host_tensors: torch.Tensor) -> None:
        """Enqueue the H2D copies of one batch (pinned tensors) on the copy stream."""
        if self._put - self._got >= self.depth:
            raise RuntimeError("DeviceFeeder: %d batches in flight, get() one first" % self.depth)
        k = self._put % self.depth
        if self._slots[k] is None:
            self._slots[k] = tuple(torch.empty(t.shape, dtype=t.dtype, device=self.device)
                                   for t in host_tensors)
            # The caching allocator may hand out a block that kernels already queued on the
            # compute stream are still using (stream-ordered reuse is only safe on THAT stream):
            # the first copy into a fresh slot has to wait for them.
            self.copy_stream.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.copy_stream):
            if self._free[k] is not None:
                self.copy_stream.wait_event(self._free[k])        # the step that used this slot ended
            for dst, src in zip(self._slots[k], host_tensors):
                dst.copy_(src, non_blocking=True)
            self._ready[k].record(self.copy_stream)
        self.bytes_per_put = sum(t.numel() * t.element_size() for t in host_tensors)
        self._put += 1

    def get(self) -> Tuple[torch.Tensor, ...]:
        """Device tensors of the oldest staged batch; the current stream waits for its copy."""
        if self._got >= self._put:
            raise RuntimeError("DeviceFeeder: nothing staged")
        k = self._got % self.depth
        torch.cuda.current_stream(self.device).wait_event(self._ready[k])
        self._cur = k
        self._got += 1
        return self._slots[k]

    def release(self) -> None:
        """Call after the work that consumed the last ``get()`` has been launched."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._free[self._cur] = ev


class ResultReader:
    """Asynchronous D2H of small per-step results (loss, metric) with a lag of ``depth-1`` steps."""

    def __init__(self, device, numel: int = 1, depth: int = 2):
        self.device = torch.device(device)
        self.depth = max(1, depth)
        self._host = [torch.zeros(numel).pin_memory() for _ in range(self.depth)]
        self._ev = [torch.cuda.Event() for _ in range(self.depth)]
        self._n = 0
        self._read = 0
        self.bytes_per_read = numel * 4

    def push(self, value: torch.Tensor) -> None:
        k = self._n % self.depth
        self._host[k].copy_(value.detach().float().reshape(-1)[:self._host[k].numel()], non_blocking=True)
        self._ev[k].record(torch.cuda.current_stream(self.device))
        self._n += 1

    def pop(self) -> Optional[torch.Tensor]:
        """Oldest unread result (blocks on its copy only); None if nothing is pending."""
        if self._read >= self._n:
            return None
        k = self._read % self.depth
        self._ev[k].synchronize()
        self._read += 1
        return self._host[k].clone()

    def pending(self) -> int:
        return self._n - self._read

"""Do the communication kernels really run WHILE a persistent GEMM holds every SM?  Times a
GEMM loop alone, a pull / push / sync loop alone, and both on two streams; overlap shows as
t(both) ~ max, serialisation as t(both) ~ sum.  One GPU (local memory as the 'ps')."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dtg_b200  # noqa: E402,F401
from dtg_b200 import ops  # noqa: E402
from dtg_b200.parallel.fabric import Fabric  # noqa: E402
from dtg_b200.parallel.peer_store import PeerSync  # noqa: E402

dev = "cuda"
M = N = K = 8192
A = torch.randn(M, K, device=dev).bfloat16(); B = torch.randn(N, K, device=dev).bfloat16()
out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
n = 53_600_272
src = torch.randn(n, device=dev); b16 = torch.empty(n, device=dev, dtype=torch.bfloat16)
g = torch.randn(n, device=dev); w = torch.zeros(n, device=dev)
sy = PeerSync(Fabric.single(), n, lr=0.1)
sy.load_master(src); sy.g32.copy_(g)
side = torch.cuda.Stream()
REPS = 6


def gemm_loop():
    for _ in range(REPS):
        ops.gemm(A, B, out)


def timed(main_fn, side_fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if side_fn is not None:
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            side_fn()
    if main_fn is not None:
        main_fn()
    if side_fn is not None:
        torch.cuda.current_stream().wait_stream(side)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


comm = {"pull_convert": lambda: [ops.pull_convert(src, b16, None) for _ in range(REPS * 4)],
        "push_axpy": lambda: [ops.push_axpy(w, g, -0.1) for _ in range(REPS * 4)],
        "sync_rs_update_ag": lambda: [sy.step() for _ in range(REPS * 2)]}
for _ in range(2):
    gemm_loop(); [f() for f in comm.values()]
tg = min(timed(gemm_loop, None) for _ in range(3))
print("GEMM loop alone            %.3f ms" % tg)
for name, fn in comm.items():
    tc = min(timed(None, fn) for _ in range(3))
    tb = min(timed(gemm_loop, fn) for _ in range(3))
    print("%-20s alone %.3f ms | with GEMM %.3f ms | sum %.3f  max %.3f  -> overlap %.0f%%" % (
        name, tc, tb, tg + tc, max(tg, tc), 100.0 * (tg + tc - tb) / min(tg, tc)))

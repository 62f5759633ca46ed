"""One bf16 GEMM shape, timed and (under ncu) captured.
usage: gemm_prof.py M N K [a_mn b_mn] [--cublas]   (--cublas: torch.matmul instead, for a side-by-side capture)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dtg_b200  # noqa: E402,F401
from dtg_b200 import ops  # noqa: E402

CUBLAS = "--cublas" in sys.argv
sys.argv = [a for a in sys.argv if a != "--cublas"]
M, N, K = (int(v) for v in sys.argv[1:4])
a_mn = len(sys.argv) > 4 and sys.argv[4] == "1"
b_mn = len(sys.argv) > 5 and sys.argv[5] == "1"
A = torch.randn((K, M) if a_mn else (M, K), device="cuda").bfloat16()
B = torch.randn((K, N) if b_mn else (N, K), device="cuda").bfloat16()
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
Af = A.t() if a_mn else A
Bf = B if b_mn else B.t()


def gemm():
    if CUBLAS:
        torch.matmul(Af, Bf, out=out)
    else:
        ops.gemm(A, B, out, a_mn=a_mn, b_mn=b_mn)


for _ in range(3):
    gemm()
torch.cuda.synchronize()
ts = []
for _ in range(10):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); gemm(); e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ts.sort()
t = ts[len(ts) // 2]
print(("cuBLAS " if CUBLAS else "") + "M=%d N=%d K=%d a_mn=%d b_mn=%d: %.3f ms  %.1f TFLOP/s" % (M, N, K, a_mn, b_mn, t, 2.0 * M * N * K / t / 1e9))
# the one launch a profiler sees (ncu --profile-from-start off)
flush.zero_()
torch.cuda.synchronize()
torch.cuda.profiler.start()
gemm()
torch.cuda.synchronize()
torch.cuda.profiler.stop()

"""4-CTA multicast GEMM: correctness against fp32 torch, then timing vs the 2-CTA kernel."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dtg_b200  # noqa: E402,F401
from dtg_b200 import ops  # noqa: E402


def rel_err(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6)).item()


torch.manual_seed(0)
for (M, N, K) in [(512, 256, 128), (1024, 768, 512), (1000, 520, 200), (2048, 1024, 4096)]:
    for a_mn, b_mn in [(False, False), (False, True), (True, True), (True, False)]:
        A = torch.randn(M, K, device="cuda").bfloat16()
        B = torch.randn(N, K, device="cuda").bfloat16()
        ref = A.float() @ B.float().t()
        a_in = A.t().contiguous() if a_mn else A
        b_in = B.t().contiguous() if b_mn else B
        out = torch.full((M, N), float("nan"), device="cuda")
        ops.gemm(a_in, b_in, out, a_mn=a_mn, b_mn=b_mn, force_bn=1024)
        torch.cuda.synchronize()
        e = rel_err(out, ref)
        print("check M=%d N=%d K=%d a_mn=%d b_mn=%d rel_err %.2e %s" % (M, N, K, a_mn, b_mn, e, "OK" if e < 1e-4 else "BAD"), flush=True)
        if not e < 1e-4:
            sys.exit(1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for (M, N, K, a_mn, b_mn) in [(8192, 8192, 8192, 0, 0), (8192, 4096, 4096, 0, 0), (8192, 4096, 4096, 0, 1),
                              (4096, 4096, 8192, 1, 1), (8192, 4096, 1024, 0, 0)]:
    A = torch.randn((K, M) if a_mn else (M, K), device="cuda").bfloat16()
    B = torch.randn((K, N) if b_mn else (N, K), device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for bn in (512, 1024):
        for _ in range(3):
            ops.gemm(A, B, out, a_mn=bool(a_mn), b_mn=bool(b_mn), force_bn=bn)
        torch.cuda.synchronize()
        ts = []
        for _ in range(8):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ops.gemm(A, B, out, a_mn=bool(a_mn), b_mn=bool(b_mn), force_bn=bn); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        t = ts[len(ts) // 2]
        print("perf M=%d N=%d K=%d a_mn=%d b_mn=%d kernel=%s: %.3f ms %.1f TFLOP/s" % (
            M, N, K, a_mn, b_mn, "4cta" if bn == 1024 else "2cta", t, 2.0 * M * N * K / t / 1e9), flush=True)

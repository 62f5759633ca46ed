"""How much of a tile's time is the exposed epilogue?  Tall and pair kernels with the epilogue
skipped (mode 100) / reduced to the TMEM read (mode 101) / complete, same shapes."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dtg_b200  # noqa
from dtg_b200 import ops
for (M, N, K) in [(8192, 8192, 8192), (8192, 4096, 4096), (8192, 4096, 1024)]:
    A = torch.randn(M, K, device="cuda").bfloat16(); B = torch.randn(N, K, device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for bn, nm in ((768, "tall"), (512, "pair")):
        res = []
        for mode in (None, 101, 100):
            f = lambda: ops.gemm(A, B, out, force_bn=bn, mode=mode)
            for _ in range(3): f()
            torch.cuda.synchronize()
            best = 1e9
            for rep in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(10): f()
                e1.record(); torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1) / 10)
            res.append(best)
        print("%dx%dx%d %s: full %.4f ms | TMEM read only %.4f | epilogue skipped %.4f  (%.0f TF full)" % (
            M, N, K, nm, res[0], res[1], res[2], 2.0 * M * N * K / res[0] / 1e9), flush=True)

import sys, torch
sys.path.insert(0, "/root/repo")
import dtg_b200  # noqa: F401
from dtg_b200 import ops
dev = "cuda"
M, N, K = 8192, 4096, 64
A = torch.randn(M, K, device=dev).bfloat16(); B = torch.randn(N, K, device=dev).bfloat16()
outb = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
for _ in range(4):
    ops.gemm(A, B, outb, force_bn=256)
torch.cuda.synchronize()

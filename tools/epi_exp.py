import sys, torch
sys.path.insert(0, "/root/repo")
import dtg_b200  # noqa: F401
from dtg_b200 import ops
dev="cuda"
M,N,K=8192,4096,4096
A=torch.randn(M,K,device=dev).bfloat16(); B=torch.randn(N,K,device=dev).bfloat16()
out=torch.empty(M,N,device=dev,dtype=torch.float32)
outb=torch.empty(M,N,device=dev,dtype=torch.bfloat16)
def t(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s,e=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(20): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e)/20
for bn in (512,256):
    print("bn",bn,"f32 store %.3f" % t(lambda: ops.gemm(A,B,out,force_bn=bn)),
      "bf16 store %.3f" % t(lambda: ops.gemm(A,B,outb,force_bn=bn)),
      "skip epi %.3f" % t(lambda: ops.gemm(A,B,out,force_bn=bn,mode=100)),
      "tmem only %.3f" % t(lambda: ops.gemm(A,B,out,force_bn=bn,mode=101)))
print("cublas %.3f" % t(lambda: torch.matmul(A,B.t())))
# small K to expose the epilogue alone
A2=A[:, :64].contiguous(); B2=B[:, :64].contiguous()
for bn in (512,256):
    print("K=64 bn",bn,"f32 %.3f"%t(lambda: ops.gemm(A2,B2,out,force_bn=bn)),"bf16 %.3f"%t(lambda: ops.gemm(A2,B2,outb,force_bn=bn)),"skip %.3f"%t(lambda: ops.gemm(A2,B2,out,force_bn=bn,mode=100)),"tmem %.3f"%t(lambda: ops.gemm(A2,B2,out,force_bn=bn,mode=101)))

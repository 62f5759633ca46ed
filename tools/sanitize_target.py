"""Small-size pass over every kernel family: the program compute-sanitizer runs
(tools/sanitize.sh) -- and, where that tool is not available, a bounds check of its own: every
GEMM output sits inside a guard band that must come back untouched.  Sizes are tiny because the tools slow kernels down 10-100x; every GEMM
kernel variant (1-CTA 64/128/256, pair, tall pair, batched), every epilogue form (LSU direct,
staged edge, TMA store, TMA reduce, coalesced red), the fused attention kernels and the whole
kernel zoo at --small sizes are touched once.  Results are compared with fp32 references so a
silent corruption is reported even when the tool finds nothing."""
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dtg_b200  # noqa: E402,F401
from dtg_b200 import ops  # noqa: E402

dev = "cuda"
torch.manual_seed(0)


def rel(o, r):
    return ((o.float() - r.float()).abs().max() / (r.float().abs().max() + 1e-6)).item()


def guarded(M, N, dtype, fill):
    """[M, N] view in the middle of a larger buffer whose border must stay untouched."""
    pad = torch.full((M + 4, N + 16), fill, device=dev, dtype=dtype)
    return pad, pad[2:M + 2, 8:N + 8]


def border_ok(pad, M, N, fill):
    chk = pad.clone()
    chk[2:M + 2, 8:N + 8] = fill
    return bool((chk == fill).all().item())


n = 0
for (M, N, K) in [(200, 136, 104), (640, 520, 264)]:
    A = torch.randn(M, K, device=dev).bfloat16()
    B = (torch.randn(N, K, device=dev) * 0.1).bfloat16()
    ref = A.float() @ B.float().t()
    for a_mn in (False, True):
        for b_mn in (False, True):
            a_in = A.t().contiguous() if a_mn else A
            b_in = B.t().contiguous() if b_mn else B
            for bn in (64, 128, 256, 512, 768):
                # outputs sit inside guard bands: a write past a tile edge shows up as a changed border
                p16, o16 = guarded(M, N, torch.bfloat16, 7.0)
                ops.gemm(a_in, b_in, o16, a_mn=a_mn, b_mn=b_mn, force_bn=bn)
                assert rel(o16, ref) < 2e-2, ("bf16", M, N, K, a_mn, b_mn, bn)
                assert border_ok(p16, M, N, 7.0), ("bf16 store wrote outside", M, N, K, a_mn, b_mn, bn)
                p32, o32 = guarded(M, N, torch.float32, 1.0)
                ops.gemm(a_in, b_in, o32, a_mn=a_mn, b_mn=b_mn, force_bn=bn, alpha=-0.5,
                         mode=ops.EPI_F32_RED_COAL)
                assert rel(o32, 1 - 0.5 * ref) < 2e-2, ("push", M, N, K, a_mn, b_mn, bn)
                assert border_ok(p32, M, N, 1.0), ("push wrote outside", M, N, K, a_mn, b_mn, bn)
                p32, o32 = guarded(M, N, torch.float32, 1.0)
                ops.gemm(a_in, b_in, o32, a_mn=a_mn, b_mn=b_mn, force_bn=bn, mode=ops.EPI_F32)
                assert rel(o32, ref) < 2e-2 and border_ok(p32, M, N, 1.0), ("f32 store", M, N, K, bn)
                n += 3
    bias = torch.randn(N, device=dev)
    mask = torch.randn(M, N, device=dev).bfloat16()
    o16 = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    pre = torch.empty_like(o16)
    ops.gemm(A, B, o16, bias=bias, act=ops.ACT_GELU, out2=pre)
    ops.gemm(A, B, o16, bias=bias, act=ops.ACT_RELU)
    ops.gemm(A, B, o16, mask=mask, mask_mode=ops.MASK_RELU)
    big = torch.zeros(M, N + 3, device=dev)
    ops.gemm(A, B, big[:, 1:N + 1], mode=ops.EPI_F32)         # unaligned: guarded scalar path
    o32 = torch.ones(M, N, device=dev)
    ops.gemm(A, B, o32, alpha=0.5, mode=ops.EPI_F32_ACC)
    n += 5
# split-K wgrad (conv-filter shape)
A = torch.randn(2048, 64, device=dev).bfloat16(); B = torch.randn(2048, 72, device=dev).bfloat16()
o = torch.empty(64, 72, device=dev)
ops.gemm(A, B, o, a_mn=True, b_mn=True)
assert rel(o, A.float().t() @ B.float()) < 2e-2
torch.cuda.synchronize()
print("sanitize_target: %d GEMM launches OK" % n, flush=True)

# fused attention forward / backward (T = 256, causal), against torch SDPA
from dtg_b200.models.transformer import tiny_transformer  # noqa: E402
from dtg_b200.models.transformer_native import NativeTransformer  # noqa: E402
tr = NativeTransformer(tiny_transformer().cfg, batch_size=2, device=dev)
tr.load_params({k: v.cuda() for k, v in tiny_transformer().init_params(seed=1).items()})
ids, labels = tr.synthetic_batch()
l0 = tr.train_step_local(ids, labels, lr=0.05)
l1 = tr.train_step_local(ids, labels, lr=0.05)
print("sanitize_target: transformer step OK (flash=%s) %.4f -> %.4f" % (tr.flash, l0, l1), flush=True)

# everything else
r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "kernel_zoo.py"), "--small"])
sys.exit(r.returncode)

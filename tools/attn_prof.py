"""Time the attention GEMMs (fused-epilogue modes and the plain batched products) in isolation.
usage: python tools/attn_prof.py [T] [B] [H] [only_mode]   (only_mode: for an ncu capture)"""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dtg_b200  # noqa: E402
from dtg_b200 import ops  # noqa: E402
from dtg_b200.ops import attn  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
H = int(sys.argv[3]) if len(sys.argv) > 3 else 16
only = sys.argv[4] if len(sys.argv) > 4 else None
hd, d = 64, H * 64
dev = "cuda"
qkv = torch.randn(B * T, 3 * d, device=dev).bfloat16()
heads = lambda t, j: t.view(B, T, 3, H, hd)[:, :, j].permute(0, 2, 1, 3)
q, k, v = (heads(qkv, j) for j in range(3))
dqkv = torch.empty_like(qkv)
dq, dk, dv = (heads(dqkv, j) for j in range(3))
P = torch.zeros(B, H, T, T, device=dev, dtype=torch.bfloat16)
dS = torch.zeros_like(P)
S32 = torch.empty(B, H, T, T, device=dev)
ctx = torch.randn(B * T, d, device=dev).bfloat16()
dctx = torch.randn(B * T, d, device=dev).bfloat16()
ch = ctx.view(B, T, H, hd).permute(0, 2, 1, 3)
dch = dctx.view(B, T, H, hd).permute(0, 2, 1, 3)
stats = torch.empty(B * H * T, (T + 127) // 128, 2, device=dev)
rm = torch.empty(B * H * T, device=dev)
ril = torch.empty(B * H * T, device=dev)
D = torch.zeros(B * H * T, device=dev)
scale = 1 / math.sqrt(hd)
s2 = scale * attn.LOG2E
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(name, fn, flops, n=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    t = ts[len(ts) // 2]
    print("%-34s %8.1f us  %7.1f TFLOP/s" % (name, t * 1e3, flops / t / 1e9), flush=True)


full = 2.0 * B * H * T * T * hd
cases = {}
for causal in (1, 0):
    f = full * (0.5 + 64.0 / T if causal else 1.0)
    tag = "causal" if causal else "full"
    cases["stats_" + tag] = (lambda c=causal: attn._call(q, k, None, (0, 0, T), T, T, hd, 5, 1.0, None, None, stats, None, s2, c), f)
    cases["probs_" + tag] = (lambda c=causal: attn._call(q, k, P.data_ptr(), (P.stride(0), P.stride(1), P.stride(2)), T, T, hd, 6, 1.0, rm, ril, None, None, s2, c), f)
    cases["dscores_" + tag] = (lambda c=causal: attn.dscores(dch, v, P, D, dS, scale, bool(c)), f)
ctx2 = torch.zeros(B * T, d, device=dev, dtype=torch.bfloat16)
for causal in (1, 0):
    f = 2 * full * (0.5 + 64.0 / T if causal else 1.0)
    tag = "causal" if causal else "full"
    cases["fused_fwd_storeP_" + tag] = (lambda c=causal: attn.fused_forward(q, k, v, ctx2, rm, ril, scale, bool(c), P=P), f)
    cases["fused_fwd_noP_" + tag] = (lambda c=causal: attn.fused_forward(q, k, v, ctx2, rm, ril, scale, bool(c)), f)
for causal in (1, 0):
    f = 4 * full * (0.5 + 64.0 / T if causal else 1.0)
    tag = "causal" if causal else "full"
    cases["fused_bwd_" + tag] = (lambda c=causal: attn.fused_backward(q, k, v, dk, dv, dch, dS, rm, ril, D, scale, bool(c)), f)
cases["rowdot"] = (lambda: attn.rowdot(dctx, ctx, D, B, T, H, hd), 0.0)
cases["plain_qk_bf16"] = (lambda: ops.gemm_batched(q, k, P), full)
cases["plain_qk_f32"] = (lambda: ops.gemm_batched(q, k, S32), full)
cases["pv"] = (lambda: ops.gemm_batched(P, v, ch, b_mn=True), full)
cases["dv=P^T dctx"] = (lambda: ops.gemm_batched(P, dch, dv, a_mn=True, b_mn=True), full)
cases["dq=dS k"] = (lambda: ops.gemm_batched(dS, k, dq, b_mn=True), full)
cases["dk=dS^T q"] = (lambda: ops.gemm_batched(dS, q, dk, a_mn=True, b_mn=True), full)
attn._call(q, k, None, (0, 0, T), T, T, hd, 5, 1.0, None, None, stats, None, s2, 1)
attn._l().dtg_attn_stats_merge(stats.data_ptr(), rm.data_ptr(), ril.data_ptr(), B * H * T, (T + 127) // 128, T, 1,
                               torch.cuda.current_stream().cuda_stream)
for name, (fn, fl) in cases.items():
    if only and name != only:
        continue
    if only:
        fn(); fn(); torch.cuda.synchronize()
    else:
        timeit(name, fn, fl)

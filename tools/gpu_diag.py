"""GPU bring-up diagnostics: every kernel family against a PyTorch fp32 reference.
Each case runs in its own subprocess under a timeout so that a trap or a hang in
one kernel neither poisons the CUDA context of the others nor stalls the box.

    python tools/gpu_diag.py            # all cases
    python tools/gpu_diag.py --case gemm_kk
"""
import argparse
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rel_err(out, ref):
    import torch
    out, ref = out.float(), ref.float()
    return ((out - ref).abs().max() / (ref.abs().max() + 1e-6)).item()


def gemm_case(a_mn, b_mn, shapes, bns=(0, 64, 128, 256, 512)):
    import torch
    from dtg_b200 import ops
    torch.manual_seed(0)
    dev = "cuda"
    for (M, N, K) in shapes:
        A = torch.randn(M, K, device=dev).bfloat16()
        B = torch.randn(N, K, device=dev).bfloat16()
        ref = A.float() @ B.float().t()
        a_in = A.t().contiguous() if a_mn else A
        b_in = B.t().contiguous() if b_mn else B
        for bn in bns:
            out = torch.full((M, N), float("nan"), device=dev, dtype=torch.float32)
            try:
                ops.gemm(a_in, b_in, out, a_mn=a_mn, b_mn=b_mn, force_bn=bn)
                torch.cuda.synchronize()
                e = rel_err(out, ref)
                print("gemm a_mn=%d b_mn=%d M=%d N=%d K=%d BN=%d rel_err=%.3e %s" % (
                    a_mn, b_mn, M, N, K, bn, e, "OK" if e < 2e-2 else "BAD"), flush=True)
            except Exception as ex:
                print("gemm a_mn=%d b_mn=%d M=%d N=%d K=%d BN=%d EXC %s" % (
                    a_mn, b_mn, M, N, K, bn, ex), flush=True)
                raise


SHAPES = [(128, 256, 64), (256, 512, 256), (200, 136, 104), (384, 1000, 784), (1024, 16, 4096),
          (16, 4096, 512), (1536, 768, 1088)]


def case_gemm_kk():
    gemm_case(False, False, SHAPES)


def case_gemm_kmn():
    gemm_case(False, True, SHAPES)


def case_gemm_mnmn():
    gemm_case(True, True, SHAPES)


def case_gemm_mnk():
    gemm_case(True, False, SHAPES)


def case_gemm_epilogue():
    import torch
    import torch.nn.functional as F
    from dtg_b200 import ops
    torch.manual_seed(1)
    dev = "cuda"
    M, N, K = 300, 520, 200
    A = torch.randn(M, K, device=dev).bfloat16()
    W = (torch.randn(N, K, device=dev) * 0.1).bfloat16()
    bias = torch.randn(N, device=dev)
    z = A.float() @ W.float().t() + bias
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    ops.gemm(A, W, out, bias=bias, act=ops.ACT_RELU)
    print("epi relu+bias rel_err=%.3e" % rel_err(out, F.relu(z)))
    pre = torch.empty_like(out)
    ops.gemm(A, W, out, bias=bias, act=ops.ACT_GELU, out2=pre)
    print("epi gelu rel_err=%.3e pre rel_err=%.3e" % (
        rel_err(out, F.gelu(z, approximate="tanh")), rel_err(pre, z)))
    res = torch.randn(M, N, device=dev).bfloat16()
    ops.gemm(A, W, out, bias=bias, residual=res)
    print("epi residual rel_err=%.3e" % rel_err(out, z + res.float()))
    mask = torch.randn(M, N, device=dev).bfloat16()
    ops.gemm(A, W, out, mask=mask, mask_mode=ops.MASK_RELU)
    print("epi relu-mask rel_err=%.3e" % rel_err(out, (z - bias) * (mask.float() > 0)))
    zz = mask.float().requires_grad_(True)
    F.gelu(zz, approximate="tanh").sum().backward()
    ops.gemm(A, W, out, mask=mask, mask_mode=ops.MASK_GELU)
    print("epi gelu-mask rel_err=%.3e" % rel_err(out, (z - bias) * zz.grad))
    o32 = torch.ones(M, N, device=dev)
    ops.gemm(A, W, o32, alpha=0.5, mode=ops.EPI_F32_ACC)
    print("epi f32 acc rel_err=%.3e" % rel_err(o32, 1 + 0.5 * (z - bias)))
    o32 = torch.ones(M, N, device=dev)
    ops.gemm(A, W, o32, alpha=-0.25, mode=ops.EPI_F32_RED)
    print("epi f32 red rel_err=%.3e" % rel_err(o32, 1 - 0.25 * (z - bias)))
    # odd pitch / unaligned -> scalar path
    big = torch.zeros(M, N + 3, device=dev)
    ops.gemm(A, W, big[:, 1:N + 1], mode=ops.EPI_F32)
    print("epi unaligned rel_err=%.3e pad=%g" % (rel_err(big[:, 1:N + 1], z - bias),
                                                 big[:, 0].abs().max().item()))
    torch.cuda.synchronize()


def case_ps_kernels():
    import torch
    from dtg_b200 import ops
    torch.manual_seed(2)
    dev = "cuda"
    n = 1_000_003
    src = torch.randn(n + 1, device=dev)[:n]          # still 16B aligned base
    src = torch.randn(n, device=dev)
    b16 = torch.empty(n, device=dev, dtype=torch.bfloat16)
    f32 = torch.empty(n, device=dev)
    ops.pull_convert(src, b16, f32)
    print("pull bf16 err=%.3e f32 eq=%s" % (rel_err(b16, src.bfloat16()), torch.equal(f32, src)))
    w = torch.randn(n, device=dev); g = torch.randn(n, device=dev); w0 = w.clone()
    ops.push_axpy(w, g, -0.1)
    print("push_axpy err=%.3e" % rel_err(w, w0 - 0.1 * g))
    w = w0.clone(); acc = torch.full((n,), 0.1, device=dev)
    ops.push_adagrad(w, acc, g, 0.01, 1.0)
    acc_ref = 0.1 + g * g
    print("push_adagrad w err=%.3e acc err=%.3e" % (
        rel_err(w, w0 - 0.01 * g / acc_ref.sqrt()), rel_err(acc, acc_ref)))
    p = w0.clone(); slot = torch.full((n,), 0.1, device=dev); win = torch.ones(n, device=dev)
    pb = torch.empty(n, device=dev, dtype=torch.bfloat16)
    ops.local_step(p, g, 0.01, p_bf16=pb, slot=slot, window=win, adagrad=True)
    pref = w0 - 0.01 * g / acc_ref.sqrt()
    print("local_step adagrad p err=%.3e win err=%.3e bf16 err=%.3e" % (
        rel_err(p, pref), rel_err(win, 1 + g), rel_err(pb, pref.bfloat16())))
    p = w0.clone()
    ops.local_step(p, g, 0.01)
    print("local_step sgd p err=%.3e" % rel_err(p, w0 - 0.01 * g))
    B, C = 1000, 10
    logits = torch.randn(B, 16, device=dev); labels = torch.randint(0, C, (B,), device=dev)
    dl = torch.empty(B, 16, device=dev, dtype=torch.bfloat16); loss = torch.zeros(1, device=dev)
    ops.softmax_xent(logits, labels, dl, loss, C, 1.0 / B)
    lg = logits[:, :C].clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lg, labels); ref.backward()
    print("xent loss err=%.3e dlogits err=%.3e pad=%g" % (
        abs(loss.item() - ref.item()) / ref.item(), rel_err(dl[:, :C], lg.grad),
        dl[:, C:].float().abs().max().item()))
    dy = torch.randn(3000, 1000, device=dev).bfloat16(); tgt = torch.zeros(1000, device=dev)
    ops.bias_grad(dy, tgt, 0.5)
    print("bias_grad err=%.3e" % rel_err(tgt, 0.5 * dy.float().sum(0)))
    ctr = torch.zeros(1, device=dev, dtype=torch.int64)
    ops.counter_add(ctr, 3); ops.counter_add(ctr, 4)
    print("counter=%d" % ctr.item())
    torch.cuda.synchronize()


def case_gemm_perf():
    import torch
    from dtg_b200 import ops
    dev = "cuda"
    for (M, N, K, a_mn, b_mn) in [(8192, 4096, 4096, False, False), (8192, 4096, 4096, False, True),
                                  (4096, 4096, 8192, True, True), (8192, 8192, 8192, False, False)]:
        A = torch.randn((K, M) if a_mn else (M, K), device=dev).bfloat16()
        B = torch.randn((K, N) if b_mn else (N, K), device=dev).bfloat16()
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        for bn in (512, 256, 128):
            for _ in range(3):
                ops.gemm(A, B, out, a_mn=a_mn, b_mn=b_mn, force_bn=bn)
            torch.cuda.synchronize()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(10):
                ops.gemm(A, B, out, a_mn=a_mn, b_mn=b_mn, force_bn=bn)
            e.record(); torch.cuda.synchronize()
            ms = s.elapsed_time(e) / 10
            print("perf M=%d N=%d K=%d a_mn=%d b_mn=%d BN=%d: %.3f ms %.1f TFLOP/s" % (
                M, N, K, a_mn, b_mn, bn, ms, 2.0 * M * N * K / ms / 1e9), flush=True)
        Af = A.t().contiguous() if a_mn else A
        Bf = B.t().contiguous() if b_mn else B
        for _ in range(3):
            torch.matmul(Af, Bf.t())
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(10):
            torch.matmul(Af, Bf.t())
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / 10
        print("perf cuBLAS M=%d N=%d K=%d: %.3f ms %.1f TFLOP/s" % (
            M, N, K, ms, 2.0 * M * N * K / ms / 1e9), flush=True)


CASES = {k[5:]: v for k, v in list(globals().items()) if k.startswith("case_")}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default=None)
    ap.add_argument("--timeout", type=int, default=120)
    ap.add_argument("--only", default=None, help="comma separated subset")
    args = ap.parse_args()
    if args.case:
        CASES[args.case]()
        return 0
    names = args.only.split(",") if args.only else list(CASES)
    bad = 0
    for name in names:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--case", name],
                               stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                               timeout=args.timeout)
            out, code = r.stdout, r.returncode
        except subprocess.TimeoutExpired as ex:
            out, code = (ex.stdout or b"").decode() if isinstance(ex.stdout, bytes) else (ex.stdout or ""), -9
        print("===== %s (exit %s, %.1fs) =====" % (name, code, time.time() - t0))
        print(out[-6000:])
        bad += code != 0
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())

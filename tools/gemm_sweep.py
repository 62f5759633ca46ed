"""A/B sweep of the 2-CTA GEMM's tunables against cuBLAS on the same box, same process.

Each configuration (environment variables read once by the library) runs in its own
subprocess; inside, this kernel and torch.matmul (cuBLAS) are timed alternately on the same
operands so both see the same clocks / power state.  Correctness of every configuration is
checked first (fp32 reference).

    python tools/gemm_sweep.py                 # default grid -> stdout + gpurun_out/gemm_sweep.txt
    python tools/gemm_sweep.py --child         # one configuration (uses the current environment)
"""
import itertools
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SHAPES = [(8192, 8192, 8192, 0, 0), (8192, 4096, 4096, 0, 0), (8192, 4096, 4096, 0, 1),
          (4096, 4096, 8192, 1, 1), (8192, 4096, 784, 0, 0), (4096, 784, 8192, 1, 1),
          (8192, 3072, 1024, 0, 0), (8192, 1024, 4096, 0, 0), (8192, 50304, 1024, 0, 0)]


def child():
    import torch
    import dtg_b200  # noqa: F401
    from dtg_b200 import ops
    dev = "cuda"
    torch.manual_seed(0)
    # correctness: odd sizes, all layouts, 2-CTA kernel forced
    for a_mn, b_mn in itertools.product((False, True), repeat=2):
        M, N, K = 2560 + 8, 776, 1096
        A = torch.randn(M, K, device=dev).bfloat16()
        B = torch.randn(N, K, device=dev).bfloat16()
        ref = A.float() @ B.float().t()
        for fbn in (512, 768):
            out = torch.full((M, N), float("nan"), device=dev)
            ops.gemm(A.t().contiguous() if a_mn else A, B.t().contiguous() if b_mn else B, out,
                     a_mn=a_mn, b_mn=b_mn, force_bn=fbn)
            err = ((out - ref).abs().max() / ref.abs().max()).item()
            assert err < 2e-2, ("pair GEMM wrong", fbn, a_mn, b_mn, err)
        # coalesced-push epilogue goes through the staging buffer
        out2 = torch.ones(M, N, device=dev)
        ops.gemm(A.t().contiguous() if a_mn else A, B.t().contiguous() if b_mn else B, out2,
                 a_mn=a_mn, b_mn=b_mn, force_bn=512, mode=ops.EPI_F32_RED_COAL, alpha=-0.5)
        err2 = ((out2 - (1 - 0.5 * ref)).abs().max() / ref.abs().max()).item()
        assert err2 < 2e-2, ("COAL epilogue wrong", a_mn, b_mn, err2)
    res = []
    for (M, N, K, a_mn, b_mn) in SHAPES:
        A = torch.randn((K, M) if a_mn else (M, K), device=dev).bfloat16()
        B = torch.randn((K, N) if b_mn else (N, K), device=dev).bfloat16()
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        Af = A.t() if a_mn else A
        Bf = B if b_mn else B.t()

        def ours():
            ops.gemm(A, B, out, a_mn=bool(a_mn), b_mn=bool(b_mn))

        def cublas():
            torch.matmul(Af, Bf, out=out)
        best = {}
        for rep in range(3):
            for name, fn in (("ours", ours), ("cublas", cublas)):
                for _ in range(3):
                    fn()
                torch.cuda.synchronize()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                for _ in range(20):
                    fn()
                e.record()
                torch.cuda.synchronize()
                ms = s.elapsed_time(e) / 20
                best[name] = min(best.get(name, 1e9), ms)
        tf = lambda ms: 2.0 * M * N * K / ms / 1e9   # noqa: E731
        res.append("%dx%dx%d a%d b%d: ours %.4f ms %.0f TF | cuBLAS %.4f ms %.0f TF | ratio %.3f" % (
            M, N, K, a_mn, b_mn, best["ours"], tf(best["ours"]), best["cublas"], tf(best["cublas"]),
            best["cublas"] / best["ours"]))
    print("\n".join(res), flush=True)


def main():
    if "--child" in sys.argv:
        return child()
    grid = [{"DTG_GEMM_TALL": "0", "DTG_STORE_TMA": "0"}, {"DTG_GEMM_TALL": "0"},
            {"DTG_GEMM_TALL_MIN_K": "0"}, {}]
    if "--full" in sys.argv:
        grid += [{"DTG_GEMM_TALL": "0", "DTG_GEMM2_STAGES": "5"}, {"DTG_GEMM_TALL": "0", "DTG_GEMM_L2HINT": "2"}]
    if "--quick" in sys.argv:
        grid = grid[:2]
    lines = []
    for cfg in grid:
        env = dict(os.environ)
        env.update(cfg)
        tag = " ".join("%s=%s" % kv for kv in cfg.items()) or "default (tall 512x256 pair tiles for K >= 6144, TMA-store epilogue)"
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child"], env=env,
                               stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=240)
            body = r.stdout.strip() if r.returncode == 0 else "FAILED rc=%d\n%s" % (r.returncode, r.stdout[-2000:])
        except subprocess.TimeoutExpired:
            body = "TIMEOUT"
        block = "== %s\n%s" % (tag, body)
        print(block, flush=True)
        lines.append(block)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "gemm_sweep.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()

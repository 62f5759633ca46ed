"""One launch of every non-GEMM kernel family at the sizes the benchmark configs use: the target
of `ncu --set full` for the per-kernel evidence under profiles/ (and of compute-sanitizer with
--small).  Prints CUDA-event times and the bytes each kernel must move, so the same run also
gives the fraction-of-HBM-roofline column of profiles/RESULTS.md.

    python tools/kernel_zoo.py            # production sizes, timed
    python tools/kernel_zoo.py --small    # tiny sizes (sanitizer runs)
    python tools/kernel_zoo.py --only init   # only the kernels whose name starts with "init"
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dtg_b200  # noqa: E402,F401
from dtg_b200 import ops  # noqa: E402
from dtg_b200.ops import conv, elastic  # noqa: E402
from dtg_b200.parallel.fabric import Fabric  # noqa: E402
from dtg_b200.parallel.peer_store import PeerSync  # noqa: E402

SMALL = "--small" in sys.argv
ONLY = sys.argv[sys.argv.index("--only") + 1] if "--only" in sys.argv else None   # name prefix
dev = "cuda"
torch.manual_seed(0)
results = []
peaks = {}
try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    pass
HBM = peaks.get("hbm_gbs", 6650.0)


def finish():
    torch.cuda.synchronize()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    tag = ("_small" if SMALL else "") + (("_" + ONLY.split()[0]) if ONLY else "")
    with open(os.path.join(ROOT, "gpurun_out", "kernel_zoo%s.txt" % tag), "w") as f:
        for r in results:
            f.write("%-34s %9.1f us %9.1f MB %8.0f GB/s %.2f\n" % r)
    print("kernel_zoo: %d kernels OK" % len(results))


def run(name, nbytes, fn, reps=5):
    if ONLY is not None and not name.startswith(ONLY):
        return
    fn()
    torch.cuda.synchronize()
    flush = torch.empty(1 if SMALL else (256 << 20), dtype=torch.uint8, device=dev)
    ts = []
    for _ in range(1 if SMALL else reps):
        flush.zero_()                                   # evict L2 between repetitions
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    ms = ts[len(ts) // 2]
    # the one launch a profiler sees (ncu --profile-from-start off)
    flush.zero_()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    fn()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    gbs = nbytes / ms / 1e6
    results.append((name, ms * 1e3, nbytes / 1e6, gbs, gbs / HBM))
    print("%-34s %9.1f us %9.1f MB %8.0f GB/s  %.2f of measured HBM copy" % (
        name, ms * 1e3, nbytes / 1e6, gbs, gbs / HBM), flush=True)


# ---- device-side initialisers (GPT-2 medium's largest tensor: wte 50304 x 1024) ------------------
from dtg_b200.ops import init as dinit  # noqa: E402
ni = 4096 if SMALL else 50304 * 1024
pi = torch.empty(ni, device=dev)
run("init normal (Philox + Box-Muller)", ni * 4, lambda: dinit.init_random_(pi, "normal", 0.0, 0.02, seed=1))
run("init truncated normal", ni * 4, lambda: dinit.init_random_(pi, "trunc_normal", 0.0, 1.0, seed=1))
run("init uniform", ni * 4, lambda: dinit.init_random_(pi, "uniform", -0.1, 0.1, seed=1))
del pi
if ONLY is not None and ONLY.startswith("init"):
    finish()
    sys.exit(0)

# ---- parameter-server data plane (wide MLP: 53.6 M parameters) --------------------------------
n = 4096 * 25 if SMALL else 53_600_272
src = torch.randn(n, device=dev); b16 = torch.empty(n, device=dev, dtype=torch.bfloat16)
f32 = torch.empty(n, device=dev); g = torch.randn(n, device=dev)
acc = torch.full((n,), 0.1, device=dev); win = torch.zeros(n, device=dev)
run("pull_convert (fp32 -> fp32+bf16)", n * 10, lambda: ops.pull_convert(src, b16, f32))
run("push_axpy (red.add v4)", n * 12, lambda: ops.push_axpy(f32, g, -0.1))
run("push_adagrad (atom acc + red w)", n * 20, lambda: ops.push_adagrad(f32, acc, g, 0.01, 1.0))
run("local_step sgd + window + bf16", n * 22, lambda: ops.local_step(f32, g, 0.01, p_bf16=b16, window=win))
run("local_step adagrad + window", n * 30,
    lambda: ops.local_step(f32, g, 0.01, p_bf16=b16, slot=acc, window=win, adagrad=True))
run("cast_f32_bf16", n * 6, lambda: ops.cast_f32_bf16(src, b16))
v = torch.zeros(n, device=dev)
run("easgd_elastic (pull+push fused)", n * 22, lambda: elastic.easgd_elastic(src, f32, b16, 0.3, velocity=v, delta=0.9))
run("momentum_step (Nesterov)", n * 26, lambda: elastic.momentum_step(f32, v, g, b16, 0.01, 0.9))
step = torch.tensor([7], device=dev); seen = torch.tensor([5], device=dev)
st = torch.zeros(1, dtype=torch.int64, device=dev)
run("push_axpy_dyn (1/staleness)", n * 12, lambda: elastic.push_axpy_dyn(f32, g, -0.5, step, seen, st))
ctr = torch.zeros(1, device=dev, dtype=torch.int64)
run("counter_add", 8, lambda: ops.counter_add(ctr, 1))
sy = PeerSync(Fabric.single(), n, lr=0.1, adagrad=False)
sy.load_master(src); sy.g32.copy_(g)
run("sync_rs_update_ag (world 1, SGD)", n * 18, lambda: sy.step())
del sy, src, f32, g, acc, win, v, b16
torch.cuda.empty_cache()

# ---- loss / bias ------------------------------------------------------------------------------
B = 256 if SMALL else 8192
logits = torch.randn(B, 16, device=dev); labels = torch.randint(0, 10, (B,), device=dev)
dl = torch.empty(B, 16, device=dev, dtype=torch.bfloat16); loss = torch.zeros(1, device=dev)
run("softmax_xent (10 classes, row16)", B * 16 * 6, lambda: ops.softmax_xent(logits, labels, dl, loss, 10, 1.0 / B))
R, V = (64, 1000) if SMALL else (8192, 50304)
lg = torch.randn(R, V, device=dev); lb = torch.randint(0, V, (R,), device=dev)
dlg = torch.empty(R, V, device=dev, dtype=torch.bfloat16)
run("softmax_xent (GPT-2 vocab, fp32 in)", R * V * 6, lambda: ops.softmax_xent(lg, lb, dlg, loss, V, 1.0 / R))
del lg, dlg
lb16 = (torch.randn(R, V, device=dev) * 3).bfloat16()
run("softmax_xent_bf16 in place (GPT-2 vocab)", R * V * 4, lambda: ops.softmax_xent_bf16_(lb16, lb, loss, V - 47, 1.0 / R))
del lb16
dy = torch.randn(B, 4096 if not SMALL else 128, device=dev).bfloat16()
tgt = torch.zeros(dy.shape[1], device=dev)
run("bias_grad (column sums)", dy.numel() * 2, lambda: ops.bias_grad(dy, tgt, 0.5))

# ---- transformer glue (GPT-2 medium: 8192 rows x 1024) -------------------------------------------
R, C = (64, 256) if SMALL else (8192, 1024)
x = torch.randn(R, C, device=dev).bfloat16(); gm = torch.randn(C, device=dev); bt = torch.randn(C, device=dev)
y = torch.empty_like(x); mean = torch.empty(R, device=dev); rstd = torch.empty(R, device=dev)
run("layernorm_fwd", R * C * 4, lambda: ops.layernorm_fwd(x, gm, bt, y, mean, rstd))
dyl = torch.randn(R, C, device=dev).bfloat16(); dx = torch.empty_like(x)
dg = torch.zeros(C, device=dev); db = torch.zeros(C, device=dev)
run("layernorm_bwd", R * C * 6, lambda: ops.layernorm_bwd(dyl, x, gm, mean, rstd, dx, dg, db, 1.0))
Vv, T = (500, 64) if SMALL else (50304, 1024)
wte = torch.randn(Vv, C, device=dev).bfloat16(); wpe = torch.randn(T, C, device=dev).bfloat16()
ids = torch.randint(0, Vv, (R,), device=dev); eo = torch.empty(R, C, device=dev, dtype=torch.bfloat16)
run("embedding_fwd", R * C * 6, lambda: ops.embedding_fwd(ids, wte, wpe, eo, T))
dwte = torch.zeros(Vv, C, device=dev); dwpe = torch.zeros(T, C, device=dev)
run("embedding_bwd", R * C * 10, lambda: ops.embedding_bwd(ids, eo, dwte, dwpe, T, 1.0))
run("add_bf16", R * C * 6, lambda: ops.add_bf16(x, y, dx))
del wte, dwte

# ---- conv-net glue (ResNet-50 stage 1: 256 x 56 x 56 x 64) -------------------------------------
N, H, W, Ci = (2, 12, 12, 16) if SMALL else (256, 56, 56, 64)
xi = torch.randn(N, H, W, Ci, device=dev).bfloat16()
col = torch.empty(N * H * W, 9 * Ci, device=dev, dtype=torch.bfloat16)
run("im2col 3x3 s1", xi.numel() * 2 + col.numel() * 2, lambda: conv.im2col(xi, col, N, H, W, Ci, 3, 3, 1, 1))
dxi = torch.empty_like(xi)
run("col2im 3x3 s1", xi.numel() * 2 + col.numel() * 2, lambda: conv.col2im(col, dxi, N, H, W, Ci, 3, 3, 1, 1))
rows, ch = N * H * W, Ci
xb = xi.view(rows, ch); sums = torch.zeros(ch, device=dev); sumsq = torch.zeros(ch, device=dev)
gb = torch.rand(ch, device=dev) + 0.5; bb = torch.randn(ch, device=dev); yb = torch.empty_like(xb)
run("bn_stats", rows * ch * 2, lambda: conv.bn_stats(xb, sums, sumsq))
run("bn_apply + relu", rows * ch * 4, lambda: conv.bn_apply(xb, sums, sumsq, gb, bb, yb, True, None))
dyb = torch.randn(rows, ch, device=dev).bfloat16(); dxb = torch.empty_like(xb)
dgb = torch.zeros(ch, device=dev); dbb = torch.zeros(ch, device=dev); scr = torch.zeros(2, ch, device=dev)
run("bn_bwd (relu)", rows * ch * 10, lambda: conv.bn_bwd(dyb, xb, yb, sums, sumsq, gb, [scr[0], scr[1]], dxb, dgb, dbb, True, 1.0, None))
Ho = conv.out_size(H, 3, 2, 1)
yp = torch.empty(N, Ho, Ho, Ci, device=dev, dtype=torch.bfloat16)
run("maxpool_fwd 3x3 s2", xi.numel() * 2 + yp.numel() * 2, lambda: conv.maxpool_fwd(xi, yp, N, H, W, Ci, 3, 2, 1))
dyp = torch.randn_like(yp)
run("maxpool_bwd 3x3 s2", xi.numel() * 4 + yp.numel() * 4, lambda: conv.maxpool_bwd(xi, yp, dyp, dxi, N, H, W, Ci, 3, 2, 1))
# implicit-GEMM 3x3 convolution (ResNet-50 stage 1: 256 x 56 x 56 x 64 -> 64), TMA im2col operand
wc = (torch.randn(Ci, 9 * Ci, device=dev) * 0.05).bfloat16()
yc = torch.empty(N * H * W, Ci, device=dev, dtype=torch.bfloat16)
flops = 2.0 * N * H * W * Ci * 9 * Ci
run("conv3x3 fwd implicit GEMM", xi.numel() * 2 + yc.numel() * 2, lambda: conv.conv_gemm_fwd(xi, wc, yc, N, H, W, Ci, 3, 1, 1))
dwc = torch.zeros(Ci, 9 * Ci, device=dev)
run("conv3x3 wgrad implicit GEMM", xi.numel() * 2 + yc.numel() * 2, lambda: conv.conv_gemm_wgrad(yc, xi, dwc, N, H, W, Ci, 3, 1, 1))
run("conv3x3 dgrad implicit GEMM", xi.numel() * 2 + yc.numel() * 2, lambda: conv.conv_gemm_dgrad(yc, wc, dxi.view(N * H * W, Ci), N, H, W, Ci, 3, 1))
print("  (conv3x3: %.1f GFLOP each)" % (flops / 1e9))
# global average pool at its real place in ResNet-50: 256 x 7 x 7 x 2048
Na, HWa, Ca = (2, 9, 64) if SMALL else (256, 49, 2048)
xa = torch.randn(Na, HWa, Ca, device=dev).bfloat16(); dxa = torch.empty_like(xa)
ya = torch.empty(Na, Ca, device=dev, dtype=torch.bfloat16)
run("avgpool_fwd 7x7x2048", xa.numel() * 2, lambda: conv.avgpool_fwd(xa, ya, Na, HWa, Ca))
run("avgpool_bwd 7x7x2048", xa.numel() * 2, lambda: conv.avgpool_bwd(ya, dxa, Na, HWa, Ca))
finish()

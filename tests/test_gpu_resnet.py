"""Conv-net kernels and the native ResNet vs PyTorch fp32 references."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel_err(out, ref):
    out, ref = out.float(), ref.float()
    return ((out - ref).abs().max() / (ref.abs().max() + 1e-6)).item()


@pytest.fixture(scope="module")
def C():
    import dtg_b200  # noqa: F401
    from dtg_b200.ops import conv
    return conv


@pytest.mark.parametrize("k,stride,pad", [(3, 1, 1), (3, 2, 1), (7, 2, 3), (1, 2, 0)])
def test_im2col_gemm_equals_conv2d_and_col2im_is_its_adjoint(C, k, stride, pad):
    from dtg_b200 import ops
    torch.manual_seed(0)
    N, H, W, Cin, Cout = 4, 20, 20, 16, 32
    x = torch.randn(N, H, W, Cin, device="cuda").bfloat16()
    w = (torch.randn(Cout, k, k, Cin, device="cuda") * 0.1).bfloat16()
    Ho, Wo = C.out_size(H, k, stride, pad), C.out_size(W, k, stride, pad)
    col = torch.empty(N * Ho * Wo, k * k * Cin, device="cuda", dtype=torch.bfloat16)
    C.im2col(x, col, N, H, W, Cin, k, k, stride, pad)
    y = torch.empty(N * Ho * Wo, Cout, device="cuda")
    ops.gemm(col, w.view(Cout, -1), y)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), None, stride, pad)
    assert rel_err(y.view(N, Ho, Wo, Cout).permute(0, 3, 1, 2), ref) < 1e-2
    # adjoint: <im2col(x), d> == <x, col2im(d)>
    d = torch.randn_like(col)
    dx = torch.empty_like(x)
    C.col2im(d, dx, N, H, W, Cin, k, k, stride, pad)
    lhs = (col.float() * d.float()).sum()
    rhs = (x.float() * dx.float()).sum()
    assert abs(lhs - rhs) < 2e-2 * abs(lhs)


@pytest.mark.parametrize("k,stride,pad,Cin,Cout,H", [(3, 1, 1, 64, 64, 20), (3, 2, 1, 128, 128, 20),
                                                      (3, 1, 1, 64, 256, 13), (1, 2, 0, 64, 128, 12),
                                                      (3, 1, 1, 256, 512, 9), (3, 1, 1, 128, 256, 60)])
def test_implicit_gemm_conv_forward_and_wgrad(C, k, stride, pad, Cin, Cout, H):
    """Convolution as an implicit GEMM: the patch matrix is gathered by TMA im2col loads
    (forward: A operand; weight gradient: B operand) and must equal conv2d / its autograd."""
    torch.manual_seed(3)
    N, W = (5, H + 3) if H < 40 else (12, H + 3)   # non-square; the big case reaches the pair kernel
    x = torch.randn(N, H, W, Cin, device="cuda").bfloat16()
    w = (torch.randn(Cout, k, k, Cin, device="cuda") * 0.05).bfloat16()
    Ho, Wo = C.out_size(H, k, stride, pad), C.out_size(W, k, stride, pad)
    rows = N * Ho * Wo
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wr = w.float().permute(0, 3, 1, 2).requires_grad_(True)
    ref = F.conv2d(xr, wr, None, stride, pad)
    y = torch.full((rows, Cout), float("nan"), device="cuda")
    C.conv_gemm_fwd(x, w.view(Cout, -1), y, N, H, W, Cin, k, stride, pad)
    assert rel_err(y.view(N, Ho, Wo, Cout).permute(0, 3, 1, 2), ref) < 1e-2
    y16 = torch.empty(rows, Cout, device="cuda", dtype=torch.bfloat16)
    C.conv_gemm_fwd(x, w.view(Cout, -1), y16, N, H, W, Cin, k, stride, pad)
    assert rel_err(y16.view(N, Ho, Wo, Cout).permute(0, 3, 1, 2), ref) < 1e-2
    dy = torch.randn(rows, Cout, device="cuda").bfloat16()
    ref.backward(dy.float().view(N, Ho, Wo, Cout).permute(0, 3, 1, 2))
    dw_ref = wr.grad.permute(0, 2, 3, 1).reshape(Cout, -1)            # [Cout, k, k, Cin]
    dw = torch.full((Cout, k * k * Cin), float("nan"), device="cuda")
    C.conv_gemm_wgrad(dy, x, dw, N, H, W, Cin, k, stride, pad)
    assert rel_err(dw, dw_ref) < 1e-2
    if stride == 1 and k > 1:
        dx = torch.full((N * H * W, Cin), float("nan"), device="cuda", dtype=torch.bfloat16)
        C.conv_gemm_dgrad(dy, w.view(Cout, -1), dx, N, H, W, Cin, k, pad)
        assert rel_err(dx.view(N, H, W, Cin).permute(0, 3, 1, 2), xr.grad) < 1e-2
    from dtg_b200 import ops
    acc = torch.ones(Cout, k * k * Cin, device="cuda")
    C.conv_gemm_wgrad(dy, x, acc, N, H, W, Cin, k, stride, pad, alpha=-0.5, mode=ops.EPI_F32_RED)
    assert rel_err(acc, 1 - 0.5 * dw_ref) < 1e-2


@pytest.mark.parametrize("relu,res", [(True, False), (False, False), (True, True)])
def test_batchnorm_fwd_bwd(C, relu, res):
    torch.manual_seed(1)
    rows, ch = 2000, 64
    x = torch.randn(rows, ch, device="cuda").bfloat16()
    g = torch.rand(ch, device="cuda") + 0.5; b = torch.randn(ch, device="cuda")
    r = torch.randn(rows, ch, device="cuda").bfloat16() if res else None
    sums = torch.zeros(ch, device="cuda"); sumsq = torch.zeros(ch, device="cuda")
    y = torch.empty_like(x)
    C.bn_stats(x, sums, sumsq)
    C.bn_apply(x, sums, sumsq, g, b, y, relu, r)
    xr = x.float().requires_grad_(True); gr = g.clone().requires_grad_(True)
    br = b.clone().requires_grad_(True)
    rr = r.float().requires_grad_(True) if res else None
    ref = F.batch_norm(xr, None, None, gr, br, True, 0.0, 1e-5)
    if res:
        ref = ref + rr
    if relu:
        ref = F.relu(ref)
    assert rel_err(y, ref) < 2e-2
    dy = torch.randn(rows, ch, device="cuda").bfloat16()
    # mask from our own bf16 output so both sides use the same ReLU pattern
    (ref * 0 + torch.where(y.float() > 0, ref, ref.detach() * 0) if relu else ref).backward(dy.float())
    dx = torch.empty_like(x); dres = torch.empty_like(x) if res else None
    dg = torch.zeros(ch, device="cuda"); db = torch.zeros(ch, device="cuda")
    scratch = torch.zeros(2, ch, device="cuda")
    C.bn_bwd(dy, x, y, sums, sumsq, g, [scratch[0], scratch[1]], dx, dg, db, relu, 1.0, dres)
    assert rel_err(dx, xr.grad) < 3e-2
    assert rel_err(dg, gr.grad) < 2e-2 and rel_err(db, br.grad) < 2e-2
    if res:
        assert rel_err(dres, rr.grad) < 2e-2


def test_pooling(C):
    torch.manual_seed(2)
    N, H, W, ch = 3, 16, 16, 32
    x = torch.randn(N, H, W, ch, device="cuda").bfloat16()
    Ho = C.out_size(H, 3, 2, 1)
    y = torch.empty(N, Ho, Ho, ch, device="cuda", dtype=torch.bfloat16)
    C.maxpool_fwd(x, y, N, H, W, ch, 3, 2, 1)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    ref = F.max_pool2d(xr, 3, 2, 1)
    assert rel_err(y.permute(0, 3, 1, 2), ref) < 1e-3
    dy = torch.randn_like(y)
    dx = torch.empty_like(x)
    C.maxpool_bwd(x, y, dy, dx, N, H, W, ch, 3, 2, 1)
    # reference with the kernel's tie rule (every pixel equal to the window max receives the
    # gradient; bf16 inputs do tie): unfold windows, compare to the max, fold back
    xp = F.pad(x.float().permute(0, 3, 1, 2), (1, 1, 1, 1), value=float("-inf"))
    win = F.unfold(xp, 3, stride=2).view(N, ch, 9, Ho * Ho)
    hit = (win == y.float().permute(0, 3, 1, 2).reshape(N, ch, 1, Ho * Ho)).float()
    contrib = hit * dy.float().permute(0, 3, 1, 2).reshape(N, ch, 1, Ho * Ho)
    refdx = F.fold(contrib.view(N, ch * 9, Ho * Ho), (H + 2, W + 2), 3, stride=2)[:, :, 1:-1, 1:-1]
    assert rel_err(dx.permute(0, 3, 1, 2), refdx) < 2e-2
    p = torch.empty(N, ch, device="cuda", dtype=torch.bfloat16)
    C.avgpool_fwd(x, p, N, H * W, ch)
    assert rel_err(p, x.float().mean((1, 2))) < 1e-2
    dxa = torch.empty_like(x)
    C.avgpool_bwd(p, dxa, N, H * W, ch)
    assert rel_err(dxa, (p.float() / (H * W))[:, None, None, :].expand(N, H, W, ch)) < 1e-2


def test_native_resnet_matches_reference():
    import dtg_b200  # noqa: F401
    from dtg_b200.models.resnet import tiny_resnet
    from dtg_b200.models.resnet_native import NativeResNet
    torch.manual_seed(0)
    ref = tiny_resnet()
    nat = NativeResNet(ref.cfg, batch_size=64, device="cuda")
    params = {k: v.cuda() for k, v in ref.init_params(seed=3).items()}
    pr = {k: (v.bfloat16().float() if v.dim() == 2 else v) for k, v in params.items()}
    x, y = nat.synthetic_batch()
    loss_ref, g_ref = ref.forward_backward(pr, (x.bfloat16().float(), y))
    loss, g = nat.forward_backward(params, (x, y))
    assert abs(loss.item() - loss_ref.item()) < 1e-2 * abs(loss_ref.item())
    # Random labels + BatchNorm make every gradient a small difference of large terms, so
    # bf16 activations/gradients show up as element-wise noise; direction is what must agree.
    cos = {k: F.cosine_similarity(g[k].float().flatten(), g_ref[k].float().flatten(), dim=0).item()
           for k in g_ref}
    bad = {k: v for k, v in cos.items() if v < 0.8}
    assert not bad, bad
    assert sum(cos.values()) / len(cos) > 0.95
    assert rel_err(g["resnet_tiny/logits/weights"], g_ref["resnet_tiny/logits/weights"]) < 5e-2


def test_native_resnet_implicit_conv_paths_match_explicit():
    """A ResNet whose 3x3 / strided convolutions take the implicit-GEMM paths (64-channel inputs)
    must produce the gradients of the same network run with explicit im2col / col2im."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = r"""
import sys, torch
sys.path.insert(0, %r)
import dtg_b200
from dtg_b200.models.resnet import ResNet, ResNetConfig
from dtg_b200.models.resnet_native import NativeResNet
torch.manual_seed(0)
ref = ResNet(ResNetConfig((2, 1), (64, 128), 64, 10, 32, "resnet_w64"))
nat = NativeResNet(ref.cfg, batch_size=16, device="cuda")
params = {k: v.cuda() for k, v in ref.init_params(seed=3).items()}
g = torch.Generator().manual_seed(1)
x, y = nat.synthetic_batch(generator=g)
loss, grads = nat.forward_backward(params, (x, y))
torch.save({"loss": loss.cpu(), "grads": {k: v.cpu() for k, v in grads.items()},
            "implicit": sum(int(u.implicit) for b in nat.blocks for u in b["u"] + ([b["sc"]] if b["sc"] else []))},
           sys.argv[1])
""" % root
    import tempfile
    outs = []
    with tempfile.TemporaryDirectory() as d:
        for flag in ("1", "0"):
            f = os.path.join(d, "g%s.pt" % flag)
            r = subprocess.run([sys.executable, "-c", script, f], stdout=subprocess.PIPE,
                               stderr=subprocess.STDOUT, text=True, timeout=300,
                               env=dict(os.environ, DTG_CONV_IMPLICIT=flag))
            assert r.returncode == 0, r.stdout[-3000:]
            outs.append(torch.load(f))
    assert outs[0]["implicit"] >= 4 and outs[1]["implicit"] == 0
    assert abs(outs[0]["loss"].item() - outs[1]["loss"].item()) < 2e-3 * abs(outs[1]["loss"].item())
    for k in outs[1]["grads"]:
        a, b = outs[0]["grads"][k].float().flatten(), outs[1]["grads"][k].float().flatten()
        assert F.cosine_similarity(a, b, dim=0).item() > 0.98, k


def test_native_resnet_trains_under_downpour_and_adag():
    from dtg_b200.models.resnet import tiny_resnet
    from dtg_b200.models.resnet_native import NativeResNet
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import GpuWorker
    for scheme in ("downpour", "adag"):
        torch.manual_seed(0)
        nat = NativeResNet(tiny_resnet().cfg, batch_size=16, device="cuda")
        w = GpuWorker(nat, SchemeConfig(scheme=scheme, lr=0.02, local_lr=0.02, update_window=2),
                      Fabric.single())
        x, y = nat.synthetic_batch()
        losses = [float(w.step(x, y).item()) for _ in range(12)]
        assert losses[-1] < losses[0], (scheme, losses)

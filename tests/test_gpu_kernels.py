"""sm_100a kernels vs plain PyTorch fp32 references (run on the B200 box)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_err(out, ref):
    out, ref = out.float(), ref.float()
    return ((out - ref).abs().max() / (ref.abs().max() + 1e-6)).item()


@pytest.fixture(scope="module")
def ops():
    import dtg_b200  # noqa: F401
    from dtg_b200 import ops as o
    o.lib()
    return o


def test_native_library_is_loaded(ops):
    maps = open("/proc/self/maps").read()
    assert "libdtg_b200.so" in maps
    assert ops.lib().dtg_version() >= 100


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("shape", [(128, 256, 64), (200, 136, 104), (384, 1000, 784),
                                   (1024, 16, 4096), (16, 4096, 512), (640, 768, 1536)])
def test_gemm_matches_fp32_reference(ops, a_mn, b_mn, shape):
    M, N, K = shape
    torch.manual_seed(0)
    A = torch.randn(M, K, device="cuda").bfloat16()
    B = torch.randn(N, K, device="cuda").bfloat16()
    ref = A.float() @ B.float().t()
    a_in = A.t().contiguous() if a_mn else A
    b_in = B.t().contiguous() if b_mn else B
    for bn in (0, 64, 128, 256, 512):        # 512 = the 2-CTA (cta_group::2) kernel
        out = torch.full((M, N), float("nan"), device="cuda")
        ops.gemm(a_in, b_in, out, a_mn=a_mn, b_mn=b_mn, force_bn=bn)
        assert rel_err(out, ref) < 1e-4, (bn, shape)


def test_gemm_epilogues(ops):
    import torch.nn.functional as F
    torch.manual_seed(1)
    M, N, K = 300, 520, 200
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    bias = torch.randn(N, device="cuda")
    z0 = A.float() @ W.float().t()
    z = z0 + bias
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    ops.gemm(A, W, out, bias=bias, act=ops.ACT_RELU)
    assert rel_err(out, F.relu(z)) < 8e-3
    pre = torch.empty_like(out)
    ops.gemm(A, W, out, bias=bias, act=ops.ACT_GELU, out2=pre)
    assert rel_err(out, F.gelu(z, approximate="tanh")) < 8e-3 and rel_err(pre, z) < 8e-3
    res = torch.randn(M, N, device="cuda").bfloat16()
    ops.gemm(A, W, out, bias=bias, residual=res)
    assert rel_err(out, z + res.float()) < 8e-3
    mask = torch.randn(M, N, device="cuda").bfloat16()
    ops.gemm(A, W, out, mask=mask, mask_mode=ops.MASK_RELU)
    assert rel_err(out, z0 * (mask.float() > 0)) < 8e-3
    zz = mask.float().requires_grad_(True)
    F.gelu(zz, approximate="tanh").sum().backward()
    ops.gemm(A, W, out, mask=mask, mask_mode=ops.MASK_GELU)
    assert rel_err(out, z0 * zz.grad) < 8e-3
    o32 = torch.ones(M, N, device="cuda")
    ops.gemm(A, W, o32, alpha=0.5, mode=ops.EPI_F32_ACC)
    assert rel_err(o32, 1 + 0.5 * z0) < 1e-5
    o32 = torch.ones(M, N, device="cuda")
    ops.gemm(A, W, o32, alpha=-0.25, mode=ops.EPI_F32_RED)       # the Hogwild push epilogue
    assert rel_err(o32, 1 - 0.25 * z0) < 1e-5
    big = torch.zeros(M, N + 3, device="cuda")
    ops.gemm(A, W, big[:, 1:N + 1], mode=ops.EPI_F32)            # unaligned -> scalar stores
    assert rel_err(big[:, 1:N + 1], z0) < 1e-5
    assert big[:, 0].abs().max() == 0 and big[:, N + 1:].abs().max() == 0


def test_gemm_rejects_bad_operands(ops):
    A = torch.randn(64, 60, device="cuda").bfloat16()       # row pitch 120 B: not TMA-legal
    B = torch.randn(64, 60, device="cuda").bfloat16()
    out = torch.empty(64, 64, device="cuda")
    with pytest.raises(ops.NativeLibraryError):
        ops.gemm(A, B, out)
    with pytest.raises(TypeError):
        ops.gemm(A.float(), B, out)


def test_push_pull_and_optimizer_kernels(ops):
    torch.manual_seed(2)
    n = 1_000_003
    src = torch.randn(n, device="cuda")
    b16 = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    f32 = torch.empty(n, device="cuda")
    ops.pull_convert(src, b16, f32)
    assert torch.equal(f32, src) and torch.equal(b16, src.bfloat16())
    w = torch.randn(n, device="cuda"); g = torch.randn(n, device="cuda"); w0 = w.clone()
    ops.push_axpy(w, g, -0.1)
    assert rel_err(w, w0 - 0.1 * g) < 1e-6
    w = w0.clone(); acc = torch.full((n,), 0.1, device="cuda")
    ops.push_adagrad(w, acc, g, 0.01, 1.0)
    acc_ref = 0.1 + g * g
    assert rel_err(acc, acc_ref) < 1e-6 and rel_err(w, w0 - 0.01 * g / acc_ref.sqrt()) < 1e-6
    p = w0.clone(); slot = torch.full((n,), 0.1, device="cuda"); win = torch.ones(n, device="cuda")
    pb = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    ops.local_step(p, g, 0.01, p_bf16=pb, slot=slot, window=win, adagrad=True)
    pref = w0 - 0.01 * g / acc_ref.sqrt()
    assert rel_err(p, pref) < 1e-6 and rel_err(win, 1 + g) < 1e-6
    assert rel_err(pb, pref.bfloat16()) < 1e-2
    p = w0.clone()
    ops.local_step(p, g, 0.01)
    assert rel_err(p, w0 - 0.01 * g) < 1e-6


def test_softmax_xent_bias_grad_counter(ops):
    torch.manual_seed(3)
    B, C = 1000, 10
    logits = torch.randn(B, 16, device="cuda")
    labels = torch.randint(0, C, (B,), device="cuda")
    dl = torch.empty(B, 16, device="cuda", dtype=torch.bfloat16)
    loss = torch.zeros(1, device="cuda")
    ops.softmax_xent(logits, labels, dl, loss, C, 1.0 / B)
    lg = logits[:, :C].clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lg, labels); ref.backward()
    assert abs(loss.item() - ref.item()) < 1e-4 * ref.item()
    assert rel_err(dl[:, :C], lg.grad) < 1e-2 and dl[:, C:].float().abs().max() == 0
    dy = torch.randn(3000, 1000, device="cuda").bfloat16(); tgt = torch.zeros(1000, device="cuda")
    ops.bias_grad(dy, tgt, 0.5)
    assert rel_err(tgt, 0.5 * dy.float().sum(0)) < 1e-5
    ctr = torch.zeros(1, device="cuda", dtype=torch.int64)
    ops.counter_add(ctr, 3); ops.counter_add(ctr, 4)
    assert ctr.item() == 7


@pytest.mark.parametrize("shape", [(64, 576, 50000), (256, 1152, 20000), (1024, 1024, 8192)])
def test_gemm_split_k_wgrad(ops, shape):
    """Small output, huge reduction (conv / attention-projection wgrad): split-K units meet by
    red.add; store and accumulate semantics must survive the split."""
    M, N, K = shape
    torch.manual_seed(2)
    A = (torch.randn(K, M, device="cuda") * 0.5).bfloat16()      # dY   [K, M]  (MN-major A)
    B = (torch.randn(K, N, device="cuda") * 0.5).bfloat16()      # act  [K, N]  (MN-major B)
    ref = A.float().t() @ B.float()
    out = torch.full((M, N), float("nan"), device="cuda")
    ops.gemm(A, B, out, a_mn=True, b_mn=True, mode=ops.EPI_F32)
    assert rel_err(out, ref) < 1e-4
    acc = torch.ones(M, N, device="cuda")
    ops.gemm(A, B, acc, a_mn=True, b_mn=True, alpha=0.5, mode=ops.EPI_F32_ACC)
    assert rel_err(acc, 1 + 0.5 * ref) < 1e-4


_PUSH_SCRIPT = r"""
import sys, torch
sys.path.insert(0, %r)
import dtg_b200
from dtg_b200 import ops
torch.manual_seed(3)
worst = 0.0
for (M, N, K, bn) in [(304, 520, 200, 0), (1536, 776, 1096, 512), (1536, 776, 1096, 768), (4096, 1024, 2048, 0), (64, 576, 8192, 0)]:
    for a_mn, b_mn in [(False, False), (True, True)]:
        A = torch.randn(M, K, device="cuda").bfloat16()
        B = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
        ref = A.float() @ B.float().t()
        a_in = A.t().contiguous() if a_mn else A
        b_in = B.t().contiguous() if b_mn else B
        for mode in (ops.EPI_F32_RED_COAL, ops.EPI_F32_RED):
            pad = torch.full((M + 2, N + 8), 7.0, device="cuda")      # guard rows / columns around the target
            tgt = pad[1:M + 1, 4:N + 4]
            ops.gemm(a_in, b_in, tgt, a_mn=a_mn, b_mn=b_mn, alpha=-0.5, mode=mode, force_bn=bn)
            ops.gemm(a_in, b_in, tgt, a_mn=a_mn, b_mn=b_mn, alpha=-0.5, mode=mode, force_bn=bn)
            torch.cuda.synchronize()
            err = ((tgt - (7.0 - ref)).abs().max() / ref.abs().max()).item()
            worst = max(worst, err)
            assert err < 1e-4, (M, N, K, a_mn, b_mn, mode, err)
            guard = pad.clone(); guard[1:M + 1, 4:N + 4] = 7.0
            assert torch.equal(guard, torch.full_like(pad, 7.0)), "push wrote outside its target"
print("PUSH_OK %%.3e" %% worst)
"""


@pytest.mark.parametrize("push_tma", ["0", "1", "2"])
def test_push_epilogues_tma_and_lsu_forms(push_tma):
    """The lock-free push of the wgrad GEMM (Hogwild/Hogwild.py:44 fused into K2) in its three
    forms: LSU red.add (0), TMA reduce-add for remote-style targets (1, default), TMA for every
    red.add output incl. split-K partials (2).  Edge tiles, both operand layouts, 1- and 2-CTA
    kernels; the bytes around the target must stay untouched."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", _PUSH_SCRIPT % root], stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=300,
                       env=dict(os.environ, DTG_PUSH_TMA=push_tma))
    assert r.returncode == 0 and "PUSH_OK" in r.stdout, r.stdout[-3000:]


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("shape", [(1304, 520, 200), (4096, 1024, 512), (2560, 776, 1096)])
def test_tall_pair_tile_gemm(ops, a_mn, b_mn, shape):
    """512 x 256 pair tiles + 256-row tail tiles (force_bn=768 selects the pair kernels; with
    M >= 512 the planner mixes big and small tiles).  Edge rows / columns, fused epilogues."""
    import torch.nn.functional as F
    torch.manual_seed(5)
    M, N, K = shape
    A = torch.randn(M, K, device="cuda").bfloat16()
    B = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    ref = A.float() @ B.float().t()
    a_in = A.t().contiguous() if a_mn else A
    b_in = B.t().contiguous() if b_mn else B
    out = torch.full((M, N), float("nan"), device="cuda")
    ops.gemm(a_in, b_in, out, a_mn=a_mn, b_mn=b_mn, force_bn=768)
    assert rel_err(out, ref) < 1e-2
    bias = torch.randn(N, device="cuda")
    o16 = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    ops.gemm(a_in, b_in, o16, a_mn=a_mn, b_mn=b_mn, force_bn=768, bias=bias, act=ops.ACT_RELU)
    assert rel_err(o16, F.relu(ref + bias)) < 1e-2
    mask = torch.randn(M, N, device="cuda").bfloat16()
    ops.gemm(a_in, b_in, o16, a_mn=a_mn, b_mn=b_mn, force_bn=768, mask=mask, mask_mode=ops.MASK_RELU)
    assert rel_err(o16, ref * (mask.float() > 0)) < 1e-2
    acc = torch.ones(M, N, device="cuda")
    ops.gemm(a_in, b_in, acc, a_mn=a_mn, b_mn=b_mn, force_bn=768, alpha=0.5, mode=ops.EPI_F32_ACC)
    assert rel_err(acc, 1 + 0.5 * ref) < 1e-2


@pytest.mark.parametrize("B,C,ld", [(64, 50257, 50304), (300, 1000, 1024), (17, 30522, 30528), (5, 65000, 65536)])
def test_softmax_xent_bf16_in_place(ops, B, C, ld):
    """LM-head loss on bf16 logits, gradient written over the logits (padding columns zeroed)."""
    import torch.nn.functional as F
    torch.manual_seed(6)
    buf = (torch.randn(B, ld, device="cuda") * 3).bfloat16()
    labels = torch.randint(0, C, (B,), device="cuda")
    ref_in = buf[:, :C].float().requires_grad_(True)
    ref = F.cross_entropy(ref_in, labels)
    ref.backward()
    loss = torch.zeros(1, device="cuda")
    ops.softmax_xent_bf16_(buf, labels, loss, C, 1.0 / B)
    torch.cuda.synchronize()
    assert abs(loss.item() - ref.item()) < 2e-3 * abs(ref.item())
    assert rel_err(buf[:, :C], ref_in.grad) < 1e-2
    assert buf[:, C:].float().abs().max().item() == 0.0


@pytest.mark.parametrize("shape", [(8192, 1024, 4096), (8192, 1024, 3072), (2048, 1024, 1024)])
def test_dgrad_gemm_with_residual_epilogue_at_model_shapes(ops, shape):
    """The post-LN transformer's residual gradient rides on the dgrad GEMM (MN-major B,
    bf16 out + bf16 residual) at BERT-large's shapes: the pair-tile kernels, not the small one."""
    torch.manual_seed(11)
    M, N, K = shape
    dy = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(K, N, device="cuda") * 0.05).bfloat16()      # [out=K, in=N]: B is MN-major
    res = torch.randn(M, N, device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    ops.gemm(dy, w, out, b_mn=True, residual=res)
    ref = dy.float() @ w.float() + res.float()
    assert rel_err(out, ref) < 8e-3


def _philox4x32_10(ctr, key):
    """Random123's Philox4x32-10 in plain Python (checked against its known-answer vectors)."""
    c, k = list(ctr), list(key)
    for _ in range(10):
        p0, p1 = 0xD2511F53 * c[0], 0xCD9E8D57 * c[2]
        c = [(p1 >> 32) ^ c[1] ^ k[0], p1 & 0xFFFFFFFF, (p0 >> 32) ^ c[3] ^ k[1], p0 & 0xFFFFFFFF]
        k = [(k[0] + 0x9E3779B9) & 0xFFFFFFFF, (k[1] + 0xBB67AE85) & 0xFFFFFFFF]
    return c


def test_device_initialisers(ops):
    """csrc/init_kernels.cu (W15 / K16): counter-based normal / truncated normal / uniform."""
    from dtg_b200.ops import init as dinit
    assert _philox4x32_10([0, 0, 0, 0], [0, 0]) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    seed, sid = 0x1234567811223344, 77
    u = torch.empty(1000, device="cuda")
    dinit.init_random_(u, dinit.UNIFORM, 0.0, 1.0, seed=seed, stream_id=sid)
    want = [(_philox4x32_10([e, 0, 0, sid], [seed & 0xFFFFFFFF, seed >> 32])[0] >> 8) / 16777216.0
            for e in range(16)]
    assert u[:16].tolist() == pytest.approx(want, abs=1e-7)          # known answer, element by element
    assert 0.0 <= float(u.min()) and float(u.max()) < 1.0
    n = 1 << 22
    z = torch.empty(n, device="cuda")
    dinit.init_random_(z, "normal", 0.5, 2.0, seed=1)
    assert abs(float(z.mean()) - 0.5) < 5e-3 and abs(float(z.std()) - 2.0) < 5e-3
    zs = (z - 0.5) / 2.0
    assert abs(float((zs ** 4).mean()) - 3.0) < 0.05 and abs(float((zs ** 3).mean())) < 0.02
    z2 = torch.empty(n, device="cuda")
    dinit.init_random_(z2, "normal", 0.5, 2.0, seed=1)
    assert torch.equal(z, z2)                                         # deterministic
    dinit.init_random_(z2, "normal", 0.5, 2.0, seed=2)
    assert not torch.equal(z, z2) and abs(float((z * z2).mean()) - 0.25) < 2e-2   # independent streams
    # the value of element e does not depend on how the tensor is cut or padded
    pad = torch.full((1024, 4100), 7.0, device="cuda")
    dinit.init_random_(pad[:, :4096], "normal", 0.5, 2.0, seed=1)
    assert torch.equal(pad[:, :4096].reshape(-1), z) and bool((pad[:, 4096:] == 7.0).all())
    t = torch.empty(n, device="cuda")
    dinit.init_random_(t, "trunc_normal", 0.0, 1.0, seed=3)           # tf.truncated_normal
    assert float(t.abs().max()) <= 2.0 and abs(float(t.std()) - 0.8796) < 3e-3
    assert abs(float(t.mean())) < 3e-3
    with pytest.raises(ValueError):
        dinit.init_random_(torch.empty(4, dtype=torch.float64), "normal")


def test_native_models_initialise_on_device():
    import dtg_b200  # noqa: F401
    from dtg_b200.models.mlp_native import NativeMLP
    from dtg_b200.models.transformer import tiny_transformer
    from dtg_b200.models.transformer_native import NativeTransformer
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import GpuWorker
    a = NativeMLP([784, 512, 10], 64, "cuda")
    b = NativeMLP([784, 512, 10], 64, "cuda", num_shards=2, sharding="balanced")
    assert a.init_on_device(seed=5) == 4 and b.init_on_device(seed=5) == 4
    pa, pb = a.export_params(), b.export_params()
    for k in pa:                                       # same seed: same bits, whatever the sharding
        assert torch.equal(pa[k], pb[k]), k
    lim = (6.0 / (784 + 512)) ** 0.5
    w = pa["dense/kernel"]
    assert float(w.abs().max()) <= lim and abs(float(w.std()) - lim / 3 ** 0.5) < 1e-3
    assert float(pa["dense/bias"].abs().max()) == 0.0
    last = a.layout.view(a.p32, "dense_1/kernel")
    assert last.shape[0] == 16 and float(last[10:].abs().max()) == 0.0      # padding stays zero
    assert rel_err(a.p16, a.p32) < 1e-2
    c = a.export_params()
    a.init_on_device(seed=6)
    assert not torch.equal(a.export_params()["dense/kernel"], c["dense/kernel"])
    t = NativeTransformer(tiny_transformer().cfg, batch_size=2, device="cuda")
    t.init_on_device(seed=1)
    pt = t.export_params()
    assert abs(float(pt["model/wte"].std()) - 0.02) < 1e-3
    assert bool((pt["model/h0/ln_1/g"] == 1).all()) and bool((pt["model/h1/mlp/c_fc/b"] == 0).all())
    assert abs(float(pt["model/h0/attn/c_proj/w"].std()) - 0.01) < 1e-3      # 0.02 / sqrt(2 * 2)
    torch.manual_seed(0)
    m = NativeMLP([784, 256, 10], 256, "cuda")
    wk = GpuWorker(m, SchemeConfig(scheme="ssgd", lr=0.05), Fabric.single(), device_init=True)
    x, y = m.synthetic_batch()
    losses = [float(wk.step(x, y).item()) for _ in range(30)]
    assert losses[-1] < losses[0] * 0.9, losses

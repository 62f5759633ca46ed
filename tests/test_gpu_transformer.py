"""Transformer kernels and the native GPT-2/BERT block vs PyTorch fp32 references."""
import math

import pytest
import torch
import torch.nn.functional as F

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


def test_layernorm_fwd_bwd(ops):
    torch.manual_seed(0)
    R, C = 1000, 1024
    x = torch.randn(R, C, device="cuda").bfloat16()
    g = torch.randn(C, device="cuda"); b = torch.randn(C, device="cuda")
    y = torch.empty_like(x); mean = torch.empty(R, device="cuda"); rstd = torch.empty(R, device="cuda")
    ops.layernorm_fwd(x, g, b, y, mean, rstd)
    xr = x.float().requires_grad_(True); gr = g.clone().requires_grad_(True); br = b.clone().requires_grad_(True)
    ref = F.layer_norm(xr, (C,), gr, br, 1e-5)
    assert rel_err(y, ref) < 1e-2
    dy = torch.randn(R, C, device="cuda").bfloat16()
    add = torch.randn(R, C, device="cuda").bfloat16()
    ref.backward(dy.float())
    dx = torch.empty_like(x); dg = torch.zeros(C, device="cuda"); db = torch.zeros(C, device="cuda")
    ops.layernorm_bwd(dy, x, g, mean, rstd, dx, dg, db, 0.5, dx_add=add)
    assert rel_err(dx, xr.grad + add.float()) < 2e-2
    assert rel_err(dg, 0.5 * gr.grad) < 1e-3 and rel_err(db, 0.5 * br.grad) < 1e-3


def test_embedding_fwd_bwd(ops):
    torch.manual_seed(1)
    V, T, D, B = 500, 64, 128, 8
    wte = torch.randn(V, D, device="cuda").bfloat16(); wpe = torch.randn(T, D, device="cuda").bfloat16()
    ids = torch.randint(0, V, (B * T,), device="cuda")
    out = torch.empty(B * T, D, device="cuda", dtype=torch.bfloat16)
    ops.embedding_fwd(ids, wte, wpe, out, T)
    ref = wte.float()[ids] + wpe.float().repeat(B, 1)
    assert rel_err(out, ref) < 1e-2
    dout = torch.randn(B * T, D, device="cuda").bfloat16()
    dwte = torch.zeros(V, D, device="cuda"); dwpe = torch.zeros(T, D, device="cuda")
    ops.embedding_bwd(ids, dout, dwte, dwpe, T, 2.0)
    rt = torch.zeros(V, D, device="cuda").index_add_(0, ids, 2.0 * dout.float())
    rp = 2.0 * dout.float().view(B, T, D).sum(0)
    assert rel_err(dwte, rt) < 1e-5 and rel_err(dwpe, rp) < 1e-5


@pytest.mark.parametrize("causal", [True, False])
def test_attention_softmax_fwd_bwd(ops, causal):
    torch.manual_seed(2)
    B, H, T = 2, 3, 256
    S = torch.randn(B, H, T, T, device="cuda") * 3
    P = torch.empty(B, H, T, T, device="cuda", dtype=torch.bfloat16)
    ops.attn_softmax_fwd(S, P, T, 0.125, causal)
    s = (S * 0.125).clone().requires_grad_(True)
    sm = s.masked_fill(torch.triu(torch.ones(T, T, dtype=torch.bool, device="cuda"), 1), float("-inf")) \
        if causal else s
    ref = torch.softmax(sm, -1)
    assert rel_err(P, ref) < 1e-2
    dP = torch.randn(B, H, T, T, device="cuda")
    dS = torch.empty_like(P)
    ops.attn_softmax_bwd(P, dP, dS, T, 0.125)
    # reference uses the same bf16-rounded probabilities
    p = P.float()
    refd = 0.125 * p * (dP - (p * dP).sum(-1, keepdim=True))
    assert rel_err(dS, refd) < 1e-2


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True)])
def test_batched_gemm_on_strided_head_views(ops, a_mn, b_mn):
    """Per-head views of a packed [B*T, 3*H*hd] QKV buffer, no copies."""
    torch.manual_seed(3)
    B, H, T, hd = 2, 4, 192, 64
    qkv = torch.randn(B * T, 3 * H * hd, device="cuda").bfloat16()
    heads = lambda j: qkv.view(B, T, 3, H, hd)[:, :, j].permute(0, 2, 1, 3)   # noqa: E731
    q, k, v = heads(0), heads(1), heads(2)
    if not a_mn and not b_mn:            # S = Q K^T
        out = torch.empty(B, H, T, T, device="cuda")
        ops.gemm_batched(q, k, out)
        ref = q.float() @ k.float().transpose(-1, -2)
    elif not a_mn and b_mn:              # O = P V into a strided [B*T, H*hd] buffer
        P = torch.randn(B, H, T, T, device="cuda").bfloat16()
        ctx = torch.zeros(B * T, H * hd, device="cuda", dtype=torch.bfloat16)
        out = ctx.view(B, T, H, hd).permute(0, 2, 1, 3)
        ops.gemm_batched(P, v, out, b_mn=True)
        ref = P.float() @ v.float()
    else:                                # dV = P^T dO
        P = torch.randn(B, H, T, T, device="cuda").bfloat16()
        out = torch.empty(B, H, T, hd, device="cuda")
        ops.gemm_batched(P, q, out, a_mn=True, b_mn=True)
        ref = P.float().transpose(-1, -2) @ q.float()
    assert rel_err(out, ref) < 1e-2


@pytest.mark.parametrize("post_ln", [False, True])
@pytest.mark.parametrize("causal", [True, False])
def test_native_transformer_matches_reference(causal, post_ln):
    """GPT-2 ordering (pre-LN) and BERT ordering (post-LN + embedding LayerNorm)."""
    import dtg_b200  # noqa: F401
    from dtg_b200.models.transformer import tiny_transformer
    from dtg_b200.models.transformer_native import NativeTransformer
    torch.manual_seed(0)
    ref = tiny_transformer(causal, post_ln=post_ln)
    nat = NativeTransformer(ref.cfg, batch_size=4, device="cuda")
    params = {k: v.cuda() for k, v in ref.init_params(seed=5).items()}
    for k in params:                      # make biases / LN params non-trivial
        if params[k].dim() == 1:
            params[k] = params[k] + 0.05 * torch.randn_like(params[k])
        else:
            params[k] = params[k] * 3
    pr = {k: (v.bfloat16().float() if v.dim() == 2 else v) for k, v in params.items()}
    ids, labels = nat.synthetic_batch()
    loss_ref, g_ref = ref.forward_backward(pr, (ids, labels))
    loss, g = nat.forward_backward(params, (ids, labels))
    assert abs(loss.item() - loss_ref.item()) < 3e-2 * abs(loss_ref.item())
    bad = {k: rel_err(g[k], g_ref[k]) for k in g_ref if rel_err(g[k], g_ref[k]) > 0.15}
    assert not bad, bad


def test_native_transformer_trains_and_runs_under_sync_engine():
    from dtg_b200.models.transformer import tiny_transformer
    from dtg_b200.models.transformer_native import NativeTransformer
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import GpuWorker
    torch.manual_seed(0)
    nat = NativeTransformer(tiny_transformer().cfg, batch_size=4, device="cuda")
    w = GpuWorker(nat, SchemeConfig(scheme="ssgd_lr", lr=1.0, worker_lrs=[0.05]), Fabric.single())
    ids, labels = nat.synthetic_batch()
    losses = [float(w.step(ids, labels).item()) for _ in range(15)]
    assert losses[-1] < losses[0] - 0.2, losses
    assert w.status() == 0 and w.global_step() == 15


def test_post_ln_transformer_trains_under_tower_averaging():
    """BASELINE config 5's scheme on BERT's block ordering."""
    from dtg_b200.models.transformer import tiny_transformer
    from dtg_b200.models.transformer_native import NativeTransformer
    from dtg_b200.parallel.engines import SchemeConfig
    from dtg_b200.parallel.fabric import Fabric
    from dtg_b200.parallel.gpu_engines import GpuWorker
    torch.manual_seed(0)
    nat = NativeTransformer(tiny_transformer(causal=False, post_ln=True).cfg, batch_size=4,
                            device="cuda")
    w = GpuWorker(nat, SchemeConfig(scheme="tower_avg", lr=0.05), Fabric.single())
    ids, labels = nat.synthetic_batch()
    losses = [float(w.step(ids, labels).item()) for _ in range(15)]
    assert losses[-1] < losses[0] - 0.2, losses
    assert w.status() == 0


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("T", [256, 200, 1024])
def test_fused_attention_epilogues(ops, causal, T):
    """softmax and its backward as epilogues of the score GEMMs vs an fp32 reference."""
    from dtg_b200.ops import attn
    torch.manual_seed(3)
    B, H, hd = 2, 3, 64
    d = H * hd
    qkv = (torch.randn(B * T, 3 * d, device="cuda") * 1.5).bfloat16()
    heads = lambda t, j: t.view(B, T, 3, H, hd)[:, :, j].permute(0, 2, 1, 3)
    q, k, v = (heads(qkv, j) for j in range(3))
    scale = 1.0 / math.sqrt(hd)
    P = torch.zeros(B, H, T, T, device="cuda", dtype=torch.bfloat16)
    stats = torch.empty(B * H * T, (T + 127) // 128, 2, device="cuda")
    rm, ril = torch.empty(B * H * T, device="cuda"), torch.empty(B * H * T, device="cuda")
    attn.softmax_probs(q, k, P, stats, rm, ril, scale, causal)
    S = torch.einsum("bhqc,bhkc->bhqk", q.float(), k.float()) * scale
    if causal:
        S = S.masked_fill(torch.triu(torch.ones(T, T, device="cuda", dtype=torch.bool), 1), -1e30)
    Pref = torch.softmax(S, -1)
    assert rel_err(P.float(), Pref) < 1e-2
    if causal:
        assert float(P.float().triu(1).abs().max()) == 0.0
    # backward: dS = scale * P * (dO V^T - rowsum(dO * O))
    ctx = torch.einsum("bhqk,bhkc->bhqc", P.float(), v.float()).permute(0, 2, 1, 3).reshape(B * T, d).bfloat16()
    dctx = torch.randn(B * T, d, device="cuda").bfloat16()
    D = torch.empty(B * H * T, device="cuda")
    attn.rowdot(dctx, ctx, D, B, T, H, hd)
    dch = dctx.view(B, T, H, hd).permute(0, 2, 1, 3)
    Dref = (dch.float() * ctx.view(B, T, H, hd).permute(0, 2, 1, 3).float()).sum(-1)
    assert rel_err(D.view(B, H, T), Dref) < 1e-3
    dS = torch.zeros_like(P)
    attn.dscores(dch, v, P, D, dS, scale, causal)
    dP = torch.einsum("bhqc,bhkc->bhqk", dch.float(), v.float())
    dSref = scale * P.float() * (dP - Dref.unsqueeze(-1))
    assert rel_err(dS.float(), dSref) < 1e-2


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("T,B,H", [(128, 1, 1), (256, 2, 3), (200, 2, 2), (1024, 2, 4)])
def test_fused_attention_forward_kernel(ops, causal, T, B, H):
    """tcgen05 two-phase attention forward (scores and P never leave the SM) vs fp32 reference."""
    from dtg_b200.ops import attn
    torch.manual_seed(11)
    hd, d = 64, H * 64
    qkv = (torch.randn(B * T, 3 * d, device="cuda") * 1.5).bfloat16()
    heads = lambda t, j: t.view(B, T, 3, H, hd)[:, :, j].permute(0, 2, 1, 3)
    q, k, v = (heads(qkv, j) for j in range(3))
    scale = 1.0 / math.sqrt(hd)
    P = torch.zeros(B, H, T, T, device="cuda", dtype=torch.bfloat16)
    ctx = torch.zeros(B * T, d, device="cuda", dtype=torch.bfloat16)
    rm, ril = torch.zeros(B * H * T, device="cuda"), torch.zeros(B * H * T, device="cuda")
    for _ in range(2):                              # second launch: barrier / TMEM state is clean
        attn.fused_forward(q, k, v, ctx, rm, ril, scale, causal, P=P)
    torch.cuda.synchronize()
    S = torch.einsum("bhqc,bhkc->bhqk", q.float(), k.float()) * scale
    if causal:
        S = S.masked_fill(torch.triu(torch.ones(T, T, device="cuda", dtype=torch.bool), 1), -1e30)
    Pref = torch.softmax(S, -1)
    assert rel_err(P.float(), Pref) < 1e-2
    Oref = torch.einsum("bhqk,bhkc->bhqc", Pref, v.float()).permute(0, 2, 1, 3).reshape(B * T, d)
    assert rel_err(ctx.float(), Oref) < 1.5e-2
    lse_ref = torch.logsumexp(S, -1).reshape(-1)
    lse = (rm - torch.log2(ril)) / attn.LOG2E
    assert float((lse - lse_ref).abs().max()) < 2e-2
    # without the P store: max-only first pass, exponentials summed once, O scaled by 1/l
    ctx2 = torch.zeros_like(ctx)
    rm2, ril2 = torch.zeros_like(rm), torch.zeros_like(ril)
    attn.fused_forward(q, k, v, ctx2, rm2, ril2, scale, causal)
    assert rel_err(ctx2.float(), Oref) < 1.5e-2
    lse2 = (rm2 - torch.log2(ril2)) / attn.LOG2E
    assert float((lse2 - lse_ref).abs().max()) < 2e-2


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("T,B,H", [(128, 1, 1), (256, 2, 3), (200, 2, 2), (1024, 2, 4)])
def test_fused_attention_backward_kernel(ops, causal, T, B, H):
    """dK, dV, dS from one tcgen05 kernel (P recomputed from row statistics) + dQ GEMM vs autograd."""
    from dtg_b200.ops import attn
    torch.manual_seed(12)
    hd, d = 64, H * 64
    qkv = (torch.randn(B * T, 3 * d, device="cuda") * 1.2).bfloat16()
    heads = lambda t, j: t.view(B, T, 3, H, hd)[:, :, j].permute(0, 2, 1, 3)
    q, k, v = (heads(qkv, j) for j in range(3))
    scale = 1.0 / math.sqrt(hd)
    ctx = torch.zeros(B * T, d, device="cuda", dtype=torch.bfloat16)
    rm, ril = torch.zeros(B * H * T, device="cuda"), torch.zeros(B * H * T, device="cuda")
    attn.fused_forward(q, k, v, ctx, rm, ril, scale, causal)
    dctx = torch.randn(B * T, d, device="cuda").bfloat16()
    dch = dctx.view(B, T, H, hd).permute(0, 2, 1, 3)
    D = torch.empty(B * H * T, device="cuda")
    attn.rowdot(dctx, ctx, D, B, T, H, hd)
    dqkv = torch.zeros_like(qkv)
    dq, dk, dv = (heads(dqkv, j) for j in range(3))
    dS = torch.zeros(B, H, T, T, device="cuda", dtype=torch.bfloat16)
    for _ in range(2):
        attn.fused_backward(q, k, v, dk, dv, dch, dS, rm, ril, D, scale, causal)
    ops.gemm_batched(dS, k, dq, b_mn=True)
    torch.cuda.synchronize()
    qf, kf, vf = (t.float().detach().requires_grad_(True) for t in (q, k, v))
    S = torch.einsum("bhqc,bhkc->bhqk", qf, kf) * scale
    if causal:
        S = S.masked_fill(torch.triu(torch.ones(T, T, device="cuda", dtype=torch.bool), 1), -1e30)
    O = torch.einsum("bhqk,bhkc->bhqc", torch.softmax(S, -1), vf)
    O.backward(dch.float())
    assert rel_err(dv.float(), vf.grad) < 2e-2
    assert rel_err(dk.float(), kf.grad) < 2e-2
    assert rel_err(dq.float(), qf.grad) < 2e-2

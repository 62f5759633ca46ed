"""Reference (plain PyTorch) model families on CPU: the oracles the sm_100a models are tested
against have to be right themselves."""
import pytest
import torch


@pytest.mark.parametrize("post_ln", [False, True])
@pytest.mark.parametrize("causal", [True, False])
def test_transformer_reference_orderings(dtg, causal, post_ln):
    from dtg_b200.models.transformer import bert_large, gpt2_medium, tiny_transformer
    m = tiny_transformer(causal=causal, post_ln=post_ln)
    p = m.init_params(seed=2)
    names = [s.name for s in m.param_specs()]
    assert ("model/ln_emb/g" in names) == post_ln and ("model/ln_f/g" in names) == (not post_ln)
    assert sum(v.numel() for v in p.values()) == 314368      # same count in both orderings
    ids, labels = m.synthetic_batch(3, generator=torch.Generator().manual_seed(0))
    loss, g = m.forward_backward(p, (ids, labels))
    assert torch.isfinite(loss) and set(g) == set(p)
    assert all(torch.isfinite(v).all() and v.abs().sum() > 0 for v in g.values())
    # causal: logits at position t do not depend on later tokens; bidirectional: they do
    ids2 = ids.clone()
    ids2[:, -1] = (ids2[:, -1] + 1) % m.cfg.vocab
    d = (m.logits(p, ids) - m.logits(p, ids2))[:, :-1].abs().max().item()
    assert (d == 0.0) if causal else (d > 0.0)
    if post_ln:
        # BERT ordering: the head reads a LayerNorm output (unit variance rows up to gamma/beta)
        q = {k: v.clone() for k, v in p.items()}
        lg = m.logits(q, ids)
        x_last = torch.linalg.lstsq(q["model/wte"].T.double().T, lg.reshape(-1, lg.shape[-1]).T.double()
                                    ).solution.T
        assert abs(float(x_last.var(-1, unbiased=False).mean()) - 1.0) < 0.05
    assert gpt2_medium().cfg.post_ln is False and bert_large(post_ln=True).cfg.post_ln is True


def test_transformer_flops_and_sizes(dtg):
    from dtg_b200.models.transformer import bert_large, gpt2_medium
    g, b = gpt2_medium(), bert_large()
    ng = sum(torch.Size(s.shape).numel() for s in g.param_specs())
    nb = sum(torch.Size(s.shape).numel() for s in b.param_specs())
    assert 350e6 < ng < 360e6 and 330e6 < nb < 340e6
    assert g.flops_per_sample() > 6 * ng * 1024 * 0.9

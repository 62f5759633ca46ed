"""Transformer family for BASELINE.json configs 4 and 5: GPT-2 medium (causal LM,
seq 1024) and BERT-large (bidirectional encoder, seq 512), both 24 layers, d=1024,
16 heads, ff 4096.  The reference repo has no such model (its only model is the
two-vector toy); these exist because the north star names them as the workloads for
synchronous SGD with per-worker learning rates and for tower averaging.

Parameter names follow the GPT-2 TF checkpoint layout (``model/h3/attn/c_attn/w`` ...).
Weights are stored ``[out, in]``.  Two block orderings, tanh-GELU, tied input/output embedding:

* pre-LN (GPT-2): ``x += attn(LN1(x)); x += mlp(LN2(x))``, final ``ln_f`` before the LM head;
* post-LN (BERT, ``cfg.post_ln``): ``x = LN_emb(emb); x = LN1(x + attn(x)); x = LN2(x + mlp(x))``,
  no final LayerNorm.  The embedding LayerNorm takes the place of ``ln_f`` in the parameter
  list (``model/ln_emb/*``), so both orderings have the same parameter count and FLOPs.

``bert_large()`` is the pre-LN block with BERT-large's dimensions, vocabulary and sequence
length (the measured config-5 rows); ``bert_large(post_ln=True)`` is BERT's own ordering.
Inputs are single-segment, so BERT's token-type row is a constant folded into the position
table; the MLM head's extra ``d x d`` transform (1 M of 335 M parameters) is left out.

This file is the plain-PyTorch reference (CPU tests, torch/cuBLAS baseline);
``transformer_native.py`` is the sm_100a implementation.
"""
from __future__ import annotations

import dataclasses
import math
from typing import Dict, List

import torch
import torch.nn.functional as F

from .base import Model, ParamSpec

__all__ = ["TransformerConfig", "Transformer", "gpt2_medium", "bert_large", "tiny_transformer"]


@dataclasses.dataclass
class TransformerConfig:
    vocab: int = 50257
    seq: int = 1024
    dim: int = 1024
    heads: int = 16
    layers: int = 24
    ff: int = 4096
    causal: bool = True
    name: str = "gpt2_medium"
    post_ln: bool = False

    @property
    def vocab_padded(self) -> int:
        return -(-self.vocab // 64) * 64

    @property
    def head_dim(self) -> int:
        return self.dim // self.heads


def _normal(std):
    return lambda shape, gen: torch.randn(shape, generator=gen) * std


def _ones(shape, gen):
    return torch.ones(shape)


def _zeros(shape, gen):
    return torch.zeros(shape)


class Transformer(Model):
    def __init__(self, cfg: TransformerConfig, compute_dtype: torch.dtype = torch.float32):
        self.cfg = cfg
        self.name = cfg.name
        self.compute_dtype = compute_dtype

    def layer_param_names(self, i: int) -> Dict[str, str]:
        p = "model/h%d/" % i
        return {"ln1_g": p + "ln_1/g", "ln1_b": p + "ln_1/b", "qkv_w": p + "attn/c_attn/w",
                "qkv_b": p + "attn/c_attn/b", "o_w": p + "attn/c_proj/w", "o_b": p + "attn/c_proj/b",
                "ln2_g": p + "ln_2/g", "ln2_b": p + "ln_2/b", "fc_w": p + "mlp/c_fc/w",
                "fc_b": p + "mlp/c_fc/b", "pr_w": p + "mlp/c_proj/w", "pr_b": p + "mlp/c_proj/b"}

    def param_specs(self) -> List[ParamSpec]:
        c = self.cfg
        d = c.dim

        def w(name, shape, std):
            return ParamSpec(name, shape, _normal(std), ("normal", std))

        def one(name):
            return ParamSpec(name, (d,), _ones, ("constant", 1.0))

        def zero(name, n=d):
            return ParamSpec(name, (n,), _zeros, ("constant", 0.0))
        out = [w("model/wte", (c.vocab, d), 0.02), w("model/wpe", (c.seq, d), 0.01)]
        pstd = 0.02 / math.sqrt(2 * c.layers)
        ln = self.norm_name()
        norm = [one(ln + "/g"), zero(ln + "/b")]
        if c.post_ln:
            # next to the embeddings: its gradient is the last one backward produces, so it
            # belongs to the bucket the synchronous engines reduce last
            out += norm
        for i in range(c.layers):
            n = self.layer_param_names(i)
            out += [one(n["ln1_g"]), zero(n["ln1_b"]),
                    w(n["qkv_w"], (3 * d, d), 0.02), zero(n["qkv_b"], 3 * d),
                    w(n["o_w"], (d, d), pstd), zero(n["o_b"]),
                    one(n["ln2_g"]), zero(n["ln2_b"]),
                    w(n["fc_w"], (c.ff, d), 0.02), zero(n["fc_b"], c.ff),
                    w(n["pr_w"], (d, c.ff), pstd), zero(n["pr_b"])]
        if not c.post_ln:
            out += norm
        return out

    def norm_name(self) -> str:
        """The one LayerNorm outside the blocks: final (pre-LN) or on the embeddings (post-LN)."""
        return "model/ln_emb" if self.cfg.post_ln else "model/ln_f"

    def logits(self, params: Dict[str, torch.Tensor], ids: torch.Tensor) -> torch.Tensor:
        c, dt = self.cfg, self.compute_dtype
        B, T = ids.shape
        P = {k: v.to(dt) for k, v in params.items()}
        x = P["model/wte"][ids] + P["model/wpe"][:T][None]
        post, ln = c.post_ln, self.norm_name()

        def norm(t, g, b):
            return F.layer_norm(t, (c.dim,), P[g], P[b], 1e-5)
        if post:
            x = norm(x, ln + "/g", ln + "/b")
        for i in range(c.layers):
            n = self.layer_param_names(i)
            h = x if post else norm(x, n["ln1_g"], n["ln1_b"])
            qkv = F.linear(h, P[n["qkv_w"]], P[n["qkv_b"]])
            q, k, v = qkv.view(B, T, 3, c.heads, c.head_dim).permute(2, 0, 3, 1, 4)
            s = (q @ k.transpose(-1, -2)).float() / math.sqrt(c.head_dim)
            if c.causal:
                s = s.masked_fill(torch.triu(torch.ones(T, T, dtype=torch.bool, device=s.device), 1),
                                  float("-inf"))
            ctx = (torch.softmax(s, -1).to(dt) @ v).transpose(1, 2).reshape(B, T, c.dim)
            x = x + F.linear(ctx, P[n["o_w"]], P[n["o_b"]])
            if post:
                x = norm(x, n["ln1_g"], n["ln1_b"])
            h = x if post else norm(x, n["ln2_g"], n["ln2_b"])
            f = F.gelu(F.linear(h, P[n["fc_w"]], P[n["fc_b"]]), approximate="tanh")
            x = x + F.linear(f, P[n["pr_w"]], P[n["pr_b"]])
            if post:
                x = norm(x, n["ln2_g"], n["ln2_b"])
        if not post:
            x = norm(x, ln + "/g", ln + "/b")
        return F.linear(x, P["model/wte"])

    def loss(self, params, batch):
        ids, labels = batch
        lg = self.logits(params, ids).float()
        return F.cross_entropy(lg.view(-1, lg.shape[-1]), labels.reshape(-1))

    def synthetic_batch(self, batch_size, device="cpu", generator=None):
        c = self.cfg
        ids = torch.randint(0, c.vocab, (batch_size, c.seq), generator=generator).to(device)
        labels = torch.randint(0, c.vocab, (batch_size, c.seq), generator=generator).to(device)
        return ids, labels

    def flops_per_sample(self) -> float:
        c = self.cfg
        per_tok = 6.0 * (c.layers * (4 * c.dim * c.dim + 2 * c.dim * c.ff) + c.vocab * c.dim)
        attn = 6.0 * c.layers * 2 * c.seq * c.dim * (0.5 if c.causal else 1.0)
        return c.seq * (per_tok + attn)


def gpt2_medium(**kw) -> Transformer:
    """GPT-2 medium, seq 1024 (BASELINE.json config 4): ~355 M parameters."""
    return Transformer(TransformerConfig(50257, 1024, 1024, 16, 24, 4096, True, "gpt2_medium"), **kw)


def bert_large(post_ln: bool = False, **kw) -> Transformer:
    """BERT-large dimensions, seq 512, bidirectional (BASELINE.json config 5): ~335 M.
    ``post_ln=True``: BERT's own block ordering with the embedding LayerNorm."""
    return Transformer(TransformerConfig(30522, 512, 1024, 16, 24, 4096, False, "bert_large",
                                         post_ln), **kw)


def tiny_transformer(causal=True, post_ln: bool = False, **kw) -> Transformer:
    return Transformer(TransformerConfig(320, 64, 128, 2, 2, 256, causal, "tiny", post_ln), **kw)

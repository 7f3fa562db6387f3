"""GPT-2 (small / medium shapes) for the transformer configurations of BASELINE.json
(configs 2-5). The reference has no transformer; this follows its conventions: Flax-style
parameter names, (in, out) Dense kernels, fp32 parameters with bf16 compute, functional
``init`` / ``apply``.

Parameter tree::

    wte (vocab, C), wpe (n_ctx, C),
    h_<i>: ln_1 {scale, bias}, attn {c_attn {kernel (C,3C), bias}, c_proj {kernel (C,C), bias}},
           ln_2 {scale, bias}, mlp {c_fc {kernel (C,4C), bias}, c_proj {kernel (4C,C), bias}},
    ln_f {scale, bias}          (the LM head is tied to wte)

Pre-LN blocks, GELU (tanh form), causal attention with head_dim 64. Bias+GELU and bias+residual
ride in the GEMM epilogues.
"""
from __future__ import annotations

import math
from typing import Dict

import torch

from .. import rng as _rng
from ..config import ConfigDict
from ..ops import functional as Fn
from .mlp import Module


def gpt2_config(name: str = "small", **over) -> ConfigDict:
    shapes = {
        "tiny": dict(n_layer=2, n_head=2, d_model=128, n_ctx=128, vocab_size=512),     # tests
        "small": dict(n_layer=12, n_head=12, d_model=768, n_ctx=1024, vocab_size=50257),
        "medium": dict(n_layer=24, n_head=16, d_model=1024, n_ctx=1024, vocab_size=50257),
    }
    cfg = ConfigDict(dict(shapes[name], dtype=torch.bfloat16, data_axis_name="data", name=name))
    for k, v in over.items():
        cfg[k] = v
    return cfg


def padded_vocab(v: int, multiple: int = 64) -> int:
    """Vocabulary rows are padded (50257 -> 50304) so the table shards evenly and tiles cleanly;
    padded rows are zero and never selected (logits over them are excluded from the loss)."""
    return (v + multiple - 1) // multiple * multiple


class GPT2(Module):
    def __init__(self, config: ConfigDict):
        self.config = config
        assert config.d_model == config.n_head * 64, "attention kernel needs head_dim 64"

    def init_params(self, key: _rng.Key, tokens: torch.Tensor, **kw) -> Dict:
        c, L, dev = self.config.d_model, self.config.n_layer, tokens.device
        g = _rng.uniform_generator(key)
        V = padded_vocab(self.config.vocab_size)

        def normal(shape, std):
            return (torch.randn(shape, generator=g) * std).to(dev)

        zeros = lambda n: torch.zeros(n, dtype=torch.float32, device=dev)
        ones = lambda n: torch.ones(n, dtype=torch.float32, device=dev)
        wte = normal((V, c), 0.02)
        wte[self.config.vocab_size:] = 0.0
        p: Dict = {"wte": wte, "wpe": normal((self.config.n_ctx, c), 0.01)}
        proj_std = 0.02 / math.sqrt(2 * L)
        for i in range(L):
            p[f"h_{i}"] = {
                "ln_1": {"scale": ones(c), "bias": zeros(c)},
                "attn": {"c_attn": {"kernel": normal((c, 3 * c), 0.02), "bias": zeros(3 * c)},
                         "c_proj": {"kernel": normal((c, c), proj_std), "bias": zeros(c)}},
                "ln_2": {"scale": ones(c), "bias": zeros(c)},
                "mlp": {"c_fc": {"kernel": normal((c, 4 * c), 0.02), "bias": zeros(4 * c)},
                        "c_proj": {"kernel": normal((4 * c, c), proj_std), "bias": zeros(c)}},
            }
        p["ln_f"] = {"scale": ones(c), "bias": zeros(c)}
        return p

    def block(self, p: Dict, x: torch.Tensor) -> torch.Tensor:
        a = Fn.layer_norm(x, p["ln_1"]["scale"], p["ln_1"]["bias"])
        qkv = Fn.dense(a, p["attn"]["c_attn"]["kernel"], p["attn"]["c_attn"]["bias"])
        att = Fn.causal_self_attention(qkv, self.config.n_head)
        x = (x.float() + Fn.dense(att, p["attn"]["c_proj"]["kernel"], p["attn"]["c_proj"]["bias"]).float()).to(torch.bfloat16)
        m = Fn.layer_norm(x, p["ln_2"]["scale"], p["ln_2"]["bias"])
        hdn = Fn.dense(m, p["mlp"]["c_fc"]["kernel"], p["mlp"]["c_fc"]["bias"], act="gelu")
        x = (x.float() + Fn.dense(hdn, p["mlp"]["c_proj"]["kernel"], p["mlp"]["c_proj"]["bias"]).float()).to(torch.bfloat16)
        return x

    def embed(self, params: Dict, tokens: torch.Tensor) -> torch.Tensor:
        return Fn.embedding(tokens, params["wte"], params["wpe"])

    def head(self, params: Dict, x: torch.Tensor) -> torch.Tensor:
        x = Fn.layer_norm(x, params["ln_f"]["scale"], params["ln_f"]["bias"])
        logits = Fn.tied_lm_head(x, params["wte"])
        return logits[..., : self.config.vocab_size]

    def forward(self, params: Dict, tokens: torch.Tensor, train: bool = False, rngs=None, **kw) -> torch.Tensor:
        x = self.embed(params, tokens)
        for i in range(self.config.n_layer):
            x = self.block(params[f"h_{i}"], x)
        return self.head(params, x).float()


def lm_loss_fn(params, apply_fn, batch, rng):
    """Next-token cross-entropy with the reference's metric protocol (sum, count)
    (data_paral.py:171-189): ``batch.inputs`` tokens [B, T], ``batch.labels`` targets [B, T]."""
    logits = apply_fn({"params": params}, batch.inputs, train=True, rngs={"dropout": rng})
    v = logits.shape[-1]
    loss, correct = Fn.softmax_cross_entropy_with_integer_labels(logits.reshape(-1, v), batch.labels.reshape(-1),
                                                                 return_correct=True)
    n = batch.labels.numel()
    return loss.mean(), {"loss": (loss.sum(), n), "accuracy": (correct, n)}

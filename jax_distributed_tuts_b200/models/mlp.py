"""Functional module system (init/apply with explicit parameter pytrees) and the reference's
MLP classifier.

``Classifier``: Dense(hidden, "input_dense") -> SiLU -> Dropout -> Dense(classes,
"output_dense") -> fp32 logits; bf16 compute, fp32 parameters, Flax names and (in, out)
kernel layout so the parameter pytree is the reference's
``{"input_dense": {"kernel", "bias"}, "output_dense": {"kernel", "bias"}}``
(data_paral.py:75-102; FSDP flavour param_sharding.py:194-224).

SiLU and dropout ride in the first GEMM's epilogue (``ops.functional.dense``), the fp32 cast in
the second's.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, Optional

import torch

from .. import rng as _rng
from ..ops import functional as Fn
from ..tree import Pytree


class Module:
    """``init(rngs, *args, **kw) -> {"params": tree}`` / ``apply({"params": tree}, *args, rngs=, **kw)``."""

    name: Optional[str] = None

    def init_params(self, key: _rng.Key, *args, **kw) -> Pytree:  # pragma: no cover - interface
        raise NotImplementedError

    def forward(self, params: Pytree, *args, rngs: Optional[Dict[str, _rng.Key]] = None, **kw):  # pragma: no cover
        raise NotImplementedError

    def init(self, rngs, *args, **kw) -> Dict[str, Pytree]:
        key = rngs["params"] if isinstance(rngs, dict) else rngs
        return {"params": self.init_params(key, *args, **kw)}

    def apply(self, variables: Dict[str, Pytree], *args, rngs: Optional[Dict[str, _rng.Key]] = None, **kw):
        return self.forward(variables["params"], *args, rngs=rngs, **kw)


def lecun_normal(key: _rng.Key, shape, device) -> torch.Tensor:
    """Flax's default Dense kernel init: truncated normal (±2σ) with std = sqrt(1/fan_in)."""
    fan_in = shape[0]
    std = math.sqrt(1.0 / fan_in) / 0.87962566103423978  # flax divides by the truncated-normal std
    g = _rng.uniform_generator(key)
    w = torch.empty(shape, dtype=torch.float32)
    torch.nn.init.trunc_normal_(w, mean=0.0, std=1.0, a=-2.0, b=2.0, generator=g)
    return (w * std).to(device)


class Dense(Module):
    """``flax.linen.Dense``: params ``kernel`` (in, out) fp32 lecun-normal, ``bias`` (out,) zeros;
    computes in ``dtype`` (bf16). ``act`` / ``dropout_rate`` are fused into the GEMM epilogue."""

    def __init__(self, features: int, dtype=torch.bfloat16, name: Optional[str] = None, act: Optional[str] = None,
                 dropout_rate: float = 0.0, out_fp32: bool = False):
        self.features, self.dtype, self.name = features, dtype, name
        self.act, self.dropout_rate, self.out_fp32 = act, dropout_rate, out_fp32

    def init_params(self, key, x, **kw):
        return {"kernel": lecun_normal(key, (x.shape[-1], self.features), x.device),
                "bias": torch.zeros(self.features, dtype=torch.float32, device=x.device)}

    def forward(self, params, x, *, train: bool = False, rngs=None, **kw):
        rate = self.dropout_rate if train else 0.0
        key = rngs["dropout"] if (rate > 0.0 and rngs is not None) else None
        return Fn.dense(x, params["kernel"], params["bias"], act=self.act, drop_rate=rate, drop_key=key,
                        out_fp32=self.out_fp32)


class Classifier(Module):
    """The reference's model (data_paral.py:75-102). ``dense_wrapper`` lets the FSDP flavour wrap
    each Dense with ``shard_module_params`` (param_sharding.py:201-205)."""

    def __init__(self, config, dense_wrapper: Optional[Callable[[Module], Module]] = None):
        self.config = config
        wrap = dense_wrapper or (lambda m: m)
        self.input_dense = wrap(Dense(config.hidden_size, dtype=config.dtype, name="input_dense", act="silu",
                                      dropout_rate=config.dropout_rate))
        self.output_dense = wrap(Dense(config.num_classes, dtype=config.dtype, name="output_dense",
                                       out_fp32=True))

    def init_params(self, key, x, train: bool = False, **kw):
        k1, k2 = _rng.split(key, 2)
        p1 = self.input_dense.init_params(k1, x)
        h = torch.empty(x.shape[:-1] + (self.config.hidden_size,), dtype=torch.bfloat16, device=x.device)
        p2 = self.output_dense.init_params(k2, h)
        return {"input_dense": p1, "output_dense": p2}

    def forward(self, params, x, train: bool = False, rngs=None, **kw):
        h = self.input_dense.forward(params["input_dense"], x, train=train, rngs=rngs)
        logits = self.output_dense.forward(params["output_dense"], h, train=train, rngs=rngs)
        return logits.float()

    def __call__(self, x, train: bool):  # bound-module call is not supported: use init/apply
        raise TypeError("use model.init(...) / model.apply({'params': params}, x, train=...)")

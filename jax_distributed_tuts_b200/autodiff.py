"""``value_and_grad`` over parameter pytrees (the ``jax.value_and_grad(loss_fn, has_aux=True)``
the reference calls at util.py:53,65), built on torch.autograd.

Leaves may be ``Partitioned`` boxes; the gradient tree has the same structure and boxes.
"""
from __future__ import annotations

from typing import Callable

import torch

from .tree import Pytree, is_partitioned, tree_leaves, tree_map


def value_and_grad(fn: Callable, has_aux: bool = False) -> Callable:
    def wrapped(params: Pytree, *args, **kw):
        def detach(p):
            v = p.value if is_partitioned(p) else p
            v = v.detach().requires_grad_(True)
            return p.replace(value=v) if is_partitioned(p) else v

        leaves_in = tree_map(detach, params, is_leaf=is_partitioned)
        with torch.enable_grad():
            out = fn(leaves_in, *args, **kw)
            loss, aux = out if has_aux else (out, None)
            flat = [p.value if is_partitioned(p) else p for p in tree_leaves(leaves_in, is_leaf=is_partitioned)]
            gflat = torch.autograd.grad(loss, flat, allow_unused=True)
        it = iter(gflat)

        def build(p):
            g = next(it)
            v = p.value if is_partitioned(p) else p
            if g is None:
                g = torch.zeros_like(v)
            return p.replace(value=g) if is_partitioned(p) else g

        grads = tree_map(build, leaves_in, is_leaf=is_partitioned)
        if aux is not None:
            aux = tree_map(lambda x: x.detach() if isinstance(x, torch.Tensor) else x, aux)
        loss = loss.detach()
        return ((loss, aux), grads) if has_aux else (loss, grads)

    return wrapped

"""AdamW with optax.adamw semantics and state layout.

Reference: ``optax.adamw(learning_rate)`` at data_paral.py:106-108 and param_sharding.py:235-237
with Optax defaults b1=0.9, b2=0.999, eps=1e-8 (outside the sqrt), weight_decay=1e-4 applied to
every leaf (no mask). ``opt_state`` has Optax's shape
``(ScaleByAdamState(count, mu, nu), EmptyState(), EmptyState())`` so a checkpoint renders as
``{"0": {"count", "mu", "nu"}, "1": {}, "2": {}}``.

``mu``/``nu`` inherit ``Partitioned`` boxes from the parameters, so a sharded parameter gets a
sharded optimizer state (param_sharding.py:232-239).

This module is the generic pytree implementation (any device, any leaf). The fused path is
``b200_adamw`` (stand-alone) and ``b200_fused_update`` (gradient all-reduce or reduce-scatter + AdamW + publish in one launch) over flat buffers.
"""
from __future__ import annotations

import dataclasses
from typing import Any, Tuple

import torch

from ..tree import Pytree, is_partitioned, register_dataclass, tree_map


@register_dataclass
@dataclasses.dataclass(frozen=True)
class ScaleByAdamState:
    count: Any
    mu: Pytree
    nu: Pytree


@register_dataclass
@dataclasses.dataclass(frozen=True)
class EmptyState:
    pass


def _val(x):
    return x.value if is_partitioned(x) else x


def _like(box, v):
    return box.replace(value=v) if is_partitioned(box) else v


@dataclasses.dataclass(frozen=True)
class AdamW:
    learning_rate: float      # or a schedule: count -> rate (optim/schedules.py)
    b1: float = 0.9
    b2: float = 0.999
    eps: float = 1e-8
    weight_decay: float = 1e-4

    def init(self, params: Pytree) -> Tuple[ScaleByAdamState, EmptyState, EmptyState]:
        zeros = lambda p: _like(p, torch.zeros_like(_val(p), dtype=torch.float32))
        mu = tree_map(zeros, params, is_leaf=is_partitioned)
        nu = tree_map(zeros, params, is_leaf=is_partitioned)
        return (ScaleByAdamState(count=0, mu=mu, nu=nu), EmptyState(), EmptyState())

    def update(self, grads: Pytree, state, params: Pytree):
        adam, e1, e2 = state
        t = adam.count + 1
        b1, b2 = self.b1, self.b2
        mu = tree_map(lambda m, g: _like(m, b1 * _val(m) + (1 - b1) * _val(g).float()), adam.mu, grads,
                      is_leaf=is_partitioned)
        nu = tree_map(lambda v, g: _like(v, b2 * _val(v) + (1 - b2) * _val(g).float() ** 2), adam.nu, grads,
                      is_leaf=is_partitioned)
        c1, c2 = 1.0 - b1 ** t, 1.0 - b2 ** t
        from .schedules import resolve
        lr = resolve(self.learning_rate, adam.count)      # a float, or an optax-style schedule count -> rate

        def upd(m, v, p):
            mh, vh = _val(m) / c1, _val(v) / c2
            return _like(p, -lr * (mh / (vh.sqrt() + self.eps) + self.weight_decay * _val(p).float()))

        updates = tree_map(upd, mu, nu, params, is_leaf=is_partitioned)
        return updates, (ScaleByAdamState(count=t, mu=mu, nu=nu), e1, e2)

    @staticmethod
    def apply_updates(params: Pytree, updates: Pytree) -> Pytree:
        return tree_map(lambda p, u: _like(p, (_val(p).float() + _val(u)).to(_val(p).dtype)), params, updates,
                        is_leaf=is_partitioned)


def adamw(learning_rate: float, b1: float = 0.9, b2: float = 0.999, eps: float = 1e-8,
          weight_decay: float = 1e-4) -> AdamW:
    return AdamW(learning_rate, b1, b2, eps, weight_decay)

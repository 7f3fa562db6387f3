"""TrainState / Batch / Metrics (reference: util.py:17-28).

``TrainState`` mirrors ``flax.training.train_state.TrainState`` plus the reference's extra
``rng`` field: ``step``, ``apply_fn`` (static), ``params``, ``tx`` (static), ``opt_state``,
``rng``; ``create`` initialises the optimizer state, ``apply_gradients`` runs ``tx.update`` and
returns the next state. On the GPU the update is performed in place on the state's buffers
(what ``donate_argnames=("state", "metrics")`` makes of the reference's jitted step,
data_paral.py:250): the returned state aliases the argument, which must not be reused.
"""
from __future__ import annotations

import dataclasses
from typing import Any, Callable, Dict, Optional, Tuple

import torch

from .rng import Key
from .tree import Pytree, register_dataclass, tree_map

Metrics = Dict[str, Tuple[Any, ...]]  # name -> (sum, count)


@register_dataclass
@dataclasses.dataclass(frozen=True)
class Batch:
    inputs: torch.Tensor
    labels: torch.Tensor

    def slice(self, start: int, end: int) -> "Batch":
        return tree_map(lambda x: x[start:end], self)

    def to(self, device, non_blocking: bool = False) -> "Batch":
        return tree_map(lambda x: x.to(device, non_blocking=non_blocking), self)


@register_dataclass(static=("apply_fn", "tx", "engine"))
@dataclasses.dataclass(frozen=True)
class TrainState:
    step: Any
    apply_fn: Callable
    params: Pytree
    tx: Any
    opt_state: Pytree
    rng: Optional[Key] = None
    engine: Any = None  # fused engine that owns the flat buffers these leaves view, if any

    @classmethod
    def create(cls, *, apply_fn: Callable, params: Pytree, tx, rng: Optional[Key] = None, **kw) -> "TrainState":
        opt_state = tx.init(params)
        return cls(step=0, apply_fn=apply_fn, params=params, tx=tx, opt_state=opt_state, rng=rng, **kw)

    def apply_gradients(self, *, grads: Pytree, **kw) -> "TrainState":
        updates, new_opt_state = self.tx.update(grads, self.opt_state, self.params)
        new_params = self.tx.apply_updates(self.params, updates)
        return dataclasses.replace(self, step=self.step + 1, params=new_params, opt_state=new_opt_state, **kw)

    def replace(self, **kw) -> "TrainState":
        return dataclasses.replace(self, **kw)

"""SGD with optax.sgd semantics and state layout (the north star names "the SGD/Adam update"; the
reference's scripts themselves only use ``optax.adamw``, data_paral.py:106-108).

``optax.sgd(learning_rate, momentum=None, nesterov=False)`` is ``trace`` (when momentum is set) followed by
``scale(-learning_rate)``; its state is ``(TraceState(trace), EmptyState())`` with momentum and
``(EmptyState(), EmptyState())`` without, so a checkpoint renders as ``{"0": {"trace": ...}, "1": {}}``; with
a schedule as learning rate the second element is ``ScaleByScheduleState(count)`` as in optax.
``trace`` inherits ``Partitioned`` boxes from the parameters like AdamW's moments do
(param_sharding.py:232-239): a sharded parameter gets a sharded momentum buffer.

Generic pytree implementation (any device, any leaf); it plugs into ``TrainState.create(tx=sgd(...))`` and
every generic-tier train step. The fused engines implement AdamW only.
"""
from __future__ import annotations

import dataclasses
from typing import Optional, Tuple

import torch

from ..tree import Pytree, is_partitioned, register_dataclass, tree_map
from .adamw import EmptyState, _like, _val


@register_dataclass
@dataclasses.dataclass(frozen=True)
class TraceState:
    trace: Pytree


@register_dataclass
@dataclasses.dataclass(frozen=True)
class ScaleByScheduleState:
    """optax keeps the step count of a scheduled learning rate in the ``scale_by_schedule`` element's state."""
    count: int


@dataclasses.dataclass(frozen=True)
class SGD:
    learning_rate: float
    momentum: Optional[float] = None
    nesterov: bool = False

    def init(self, params: Pytree) -> Tuple:
        lr_state = ScaleByScheduleState(count=0) if callable(self.learning_rate) else EmptyState()
        if self.momentum is None:
            return (EmptyState(), lr_state)
        zeros = lambda p: _like(p, torch.zeros_like(_val(p), dtype=torch.float32))
        return (TraceState(trace=tree_map(zeros, params, is_leaf=is_partitioned)), lr_state)

    def update(self, grads: Pytree, state, params: Pytree = None):
        from .schedules import resolve
        first, lr_state = state
        count = lr_state.count if isinstance(lr_state, ScaleByScheduleState) else 0
        lr = resolve(self.learning_rate, count)
        if isinstance(lr_state, ScaleByScheduleState):
            lr_state = ScaleByScheduleState(count=count + 1)
        if self.momentum is None:
            return tree_map(lambda g: _like(g, -lr * _val(g).float()), grads, is_leaf=is_partitioned), (first, lr_state)
        mom = self.momentum
        tr, e = first, lr_state
        new_trace = tree_map(lambda t, g: _like(t, _val(g).float() + mom * _val(t)), tr.trace, grads, is_leaf=is_partitioned)
        if self.nesterov:
            upd = tree_map(lambda t, g: _like(t, -lr * (_val(g).float() + mom * _val(t))), new_trace, grads,
                           is_leaf=is_partitioned)
        else:
            upd = tree_map(lambda t: _like(t, -lr * _val(t)), new_trace, is_leaf=is_partitioned)
        return upd, (TraceState(trace=new_trace), e)

    @staticmethod
    def apply_updates(params: Pytree, updates: Pytree) -> Pytree:
        return tree_map(lambda p, u: _like(p, (_val(p).float() + _val(u)).to(_val(p).dtype)), params, updates,
                        is_leaf=is_partitioned)


def sgd(learning_rate: float, momentum: Optional[float] = None, nesterov: bool = False) -> SGD:
    return SGD(learning_rate, momentum, nesterov)

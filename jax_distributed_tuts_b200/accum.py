"""Gradient accumulation over minibatches (reference: util.py:41-167).

``accum_grads_loop`` is the reference's live path: a Python loop over ``n_minbatch`` equal
slices, per-minibatch RNG from ``split(key, n)``, gradients *averaged* (``/ n_minbatch``,
util.py:77) and metrics *summed* (util.py:75).

``accum_grads_scan`` is the reference's ``lax.scan`` variant (one compiled body instead of n
unrolled copies, util.py:81-137): zero-initialised carries with the shapes of one minibatch
step (util.py:123-129) and the same body applied to ``narrow``-ed windows of the batch
(the ``dynamic_slice_in_dim`` of util.py:97-105). There is no tracing compiler here, so in the
generic tier the body simply runs n times; the compiled-once counterpart lives in the fused
engines, whose whole step is one captured CUDA graph (and which can fold the minibatch loop
into the GEMMs themselves, ``engine.MLPEngine(fuse_minibatches=True)``).
"""
from __future__ import annotations

from typing import Callable, Tuple

import torch

from . import rng as _rng
from .autodiff import value_and_grad
from .state import Batch, Metrics, TrainState
from .tree import Pytree, is_partitioned, tree_map


def _add(a, b):
    if is_partitioned(a):
        return a.replace(value=a.value + b.value)
    return a + b


def accum_grads_loop(batch: Batch, state: TrainState, key: _rng.Key, n_minbatch: int,
                     loss_fn: Callable) -> Tuple[Pytree, Metrics]:
    bs = batch.inputs.shape[0]
    min_batch_size = bs // n_minbatch
    keys = _rng.split(key, n_minbatch)
    grad_fn = value_and_grad(loss_fn, has_aux=True)
    grads = None
    metrics: Metrics = None
    for batch_idx in range(n_minbatch):
        start = batch_idx * min_batch_size
        minibatch = batch.slice(start, start + min_batch_size)
        (_, step_metric), step_grad = grad_fn(state.params, state.apply_fn, minibatch, keys[batch_idx])
        if grads is None:
            grads, metrics = step_grad, step_metric
        else:
            grads = tree_map(_add, grads, step_grad, is_leaf=is_partitioned)
            metrics = tree_map(lambda a, b: a + b, metrics, step_metric)
    grads = tree_map(lambda g: g.replace(value=g.value / n_minbatch) if is_partitioned(g) else g / n_minbatch,
                     grads, is_leaf=is_partitioned)
    return grads, metrics


def accum_grads_scan(batch: Batch, state: TrainState, key: _rng.Key, n_minbatch: int,
                     loss_fn: Callable) -> Tuple[Pytree, Metrics]:
    bs = batch.inputs.shape[0]
    min_batch_size = bs // n_minbatch
    keys = _rng.split(key, n_minbatch)
    grad_fn = value_and_grad(loss_fn, has_aux=True)

    def minibatch_step(idx: int):
        mb = tree_map(lambda x: x.narrow(0, idx * min_batch_size, min_batch_size), batch)
        (_, step_metric), step_grad = grad_fn(state.params, state.apply_fn, mb, keys[idx])
        return step_grad, step_metric

    # zero-initialised carry with the shapes of one step (util.py:123-129 uses eval_shape)
    g0, m0 = minibatch_step(0)
    zeros = lambda x: (x.replace(value=torch.zeros_like(x.value)) if is_partitioned(x)
                       else torch.zeros_like(x) if isinstance(x, torch.Tensor) else type(x)(0))
    grads = tree_map(zeros, g0, is_leaf=is_partitioned)
    metrics = tree_map(zeros, m0)
    carry = (grads, metrics)
    for idx in range(n_minbatch):
        sg, sm = (g0, m0) if idx == 0 else minibatch_step(idx)
        carry = (tree_map(_add, carry[0], sg, is_leaf=is_partitioned), tree_map(lambda a, b: a + b, carry[1], sm))
    grads, metrics = carry
    grads = tree_map(lambda g: g.replace(value=g.value / n_minbatch) if is_partitioned(g) else g / n_minbatch,
                     grads, is_leaf=is_partitioned)
    return grads, metrics


def accum_grads(state: TrainState, batch: Batch, key: _rng.Key, num_minibatches: int, loss_fn: Callable,
                use_scan: bool = False) -> Tuple[Pytree, Metrics]:
    """Dispatcher with the reference's signature (util.py:140-167)."""
    if use_scan:
        return accum_grads_scan(batch=batch, state=state, key=key, n_minbatch=num_minibatches, loss_fn=loss_fn)
    return accum_grads_loop(batch=batch, state=state, key=key, n_minbatch=num_minibatches, loss_fn=loss_fn)

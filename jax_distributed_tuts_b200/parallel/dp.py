"""Data parallelism: replicated parameters, per-device gradient accumulation, gradient
all-reduce-mean, replicated AdamW (reference: data_paral.py:28-34,128-251).

This file is the generic tier: it works on any mesh/backend (gloo on CPU, NCCL on GPUs) and
with any ``loss_fn``. ``make_train_step_dp`` is the analogue of
``jax.jit(shard_map(train_step_dp, ...), donate_argnames=("state", "metrics"))``
(data_paral.py:241-251); when the state was created by the fused engine
(``engine.MLPEngine`` / ``engine.GPT2Engine`` with ``mode="dp"``) the returned callable runs the CUDA-graph-captured, fused-collective
step instead.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch

from .. import rng as _rng
from ..accum import accum_grads
from ..comm import collectives as C
from ..mesh import Mesh, current_mesh
from ..ops import functional as Fn
from ..optim import adamw
from ..state import Batch, Metrics, TrainState
from ..tree import tree_map
from ..utils.profiling import named_scope


def fold_rng_over_axis(rng: _rng.Key, axis_name: str, mesh: Optional[Mesh] = None) -> _rng.Key:
    """Distinct RNG stream per coordinate of ``axis_name`` (data_paral.py:28-34)."""
    return _rng.fold_in(rng, C.axis_index(axis_name, mesh))


def init_dp(rng: _rng.Key, x: torch.Tensor, model, optimizer=None, learning_rate: float = 1e-3) -> TrainState:
    """Per-device initialisation (data_paral.py:128-147). Every rank uses the same ``rng`` so
    the parameters start replicated without communication."""
    init_rng, rng = _rng.split(rng, 2)
    variables = model.init({"params": init_rng}, x, train=False)
    params = variables.pop("params")
    tx = optimizer if optimizer is not None else adamw(learning_rate)
    # the stored key is always in rebased form (salt folded into the 64-bit state), the same
    # form every train step leaves it in
    return TrainState.create(apply_fn=model.apply, params=params, tx=tx, rng=_rng.rebase(rng))


def make_loss_fn(data_axis_name: str = "data") -> Callable:
    def loss_fn(params, apply_fn, batch: Batch, rng: _rng.Key):
        """Softmax cross-entropy with integer labels + accuracy (data_paral.py:171-189).
        Returns (mean loss, {"loss": (sum, n), "accuracy": (n_correct, n)})."""
        dropout_rng = fold_rng_over_axis(rng, data_axis_name)
        logits = apply_fn({"params": params}, batch.inputs, train=True, rngs={"dropout": dropout_rng})
        loss, correct = Fn.softmax_cross_entropy_with_integer_labels(logits, batch.labels, return_correct=True)
        bs = batch.inputs.shape[0]
        step_metrics = {"loss": (loss.sum(), bs), "accuracy": (correct, bs)}
        return loss.mean(), step_metrics

    return loss_fn


loss_fn = make_loss_fn("data")


def train_step_dp(state: TrainState, metrics: Optional[Metrics], batch: Batch, *, num_minibatches: int = 4,
                  data_axis_name: str = "data", loss_fn: Callable = loss_fn,
                  use_scan: bool = False) -> Tuple[TrainState, Metrics]:
    """One data-parallel step on this rank's slice of the batch (data_paral.py:193-238):
    split rng -> accumulate grads over minibatches -> pmean grads ("sync_grads") ->
    apply_gradients -> psum metrics ("sync_metrics") -> add to the running metrics."""
    rng, step_rng = _rng.split(state.rng, 2)
    grads, step_metrics = accum_grads(state, batch, step_rng, num_minibatches, loss_fn=loss_fn, use_scan=use_scan)
    with named_scope("sync_grads"):
        grads = tree_map(lambda g: C.pmean(g, axis_name=data_axis_name), grads)
    new_state = state.apply_gradients(grads=grads, rng=_rng.rebase(rng))
    with named_scope("sync_metrics"):
        step_metrics = tree_map(lambda x: C.psum(x, axis_name=data_axis_name), step_metrics)
    if metrics is None:
        metrics = step_metrics
    else:
        metrics = tree_map(lambda a, b: a + b, metrics, step_metrics)
    return new_state, metrics


def make_train_step_dp(mesh: Optional[Mesh] = None, *, num_minibatches: int = 4, data_axis_name: str = "data",
                       loss_fn: Callable = loss_fn, use_scan: bool = False) -> Callable:
    """Binds mesh and hyper-parameters: ``step(state, metrics, batch) -> (state, metrics)``.

    ``batch`` is this rank's shard (what ``in_specs=P("data")`` hands each device in the
    reference). States created by a fused engine dispatch to the engine's captured step."""
    mesh = mesh or current_mesh()

    def step(state: TrainState, metrics: Optional[Metrics], batch: Batch):
        if state.engine is not None:
            return state.engine.train_step(state, metrics, batch)
        with mesh:
            return train_step_dp(state, metrics, batch, num_minibatches=num_minibatches,
                                 data_axis_name=data_axis_name, loss_fn=loss_fn, use_scan=use_scan)

    return step


def shard_batch(batch: Batch, axis_name: str = "data", mesh: Optional[Mesh] = None) -> Batch:
    """This rank's rows of a global batch: the ``P("data")`` in_spec (data_paral.py:245)."""
    m = mesh or current_mesh()
    n, i = m.axis_size(axis_name), m.axis_index(axis_name)
    per = batch.inputs.shape[0] // n
    return batch.slice(i * per, (i + 1) * per)

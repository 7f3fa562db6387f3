"""Parameter sharding (FSDP / ZeRO-3 style) — reference: param_sharding.py:58-379.

Every large parameter is stored as a 1/N slice along its largest divisible dimension
(``shard_params``), gathered right before use (``gather_params``) with a custom backward that
reduce-scatters the gradient and divides by the axis size (``gather_arr_mean_grads``), so
gradients and AdamW state stay sharded.

Generic tier: the gather is a collective call followed by the consuming op. The fused tier
(``engine.MLPEngine`` / ``engine.GPT2Engine`` with ``mode="fsdp"``) never materialises the gathered weight: the consuming tcgen05 GEMM's
TMA producer reads the peer shards in place, and the reduce-scatter is fused with the sharded
AdamW step.
"""
from __future__ import annotations

import logging
from typing import Callable, Optional, Sequence, Tuple

import numpy as np
import torch

from .. import rng as _rng
from ..accum import accum_grads
from ..comm import collectives as C
from ..mesh import Mesh, current_mesh
from ..models.mlp import Module
from ..optim import adamw
from ..state import Batch, Metrics, TrainState
from ..tree import Partitioned, Pytree, is_partitioned, tree_map
from ..utils.profiling import named_scope
from .dp import make_loss_fn

log = logging.getLogger("b200dist.fsdp")


def choose_shard_dim(shape: Sequence[int], names: Sequence[Optional[str]], axis_size: int) -> Optional[int]:
    """The reference's rule (param_sharding.py:94-113): dims in descending size order, first one
    divisible by the axis size and not already named. ``None`` if there is none."""
    for i in np.argsort(shape)[::-1]:
        i = int(i)
        if shape[i] % axis_size == 0 and names[i] is None:
            return i
    return None


@named_scope("shard_params")
def shard_params(params: Pytree, axis_name: str, min_weight_size: int = 2 ** 18,
                 mesh: Optional[Mesh] = None) -> Pytree:
    """Slices every leaf that is large enough along its largest divisible dim and boxes it as
    ``Partitioned(value=<local slice>, names=(..., axis_name, ...))`` (param_sharding.py:58-125)."""
    m = mesh or current_mesh()
    axis_idx = m.axis_index(axis_name)
    axis_size = m.axis_size(axis_name)

    def _split(x):
        if is_partitioned(x):
            value, names = x.value, x.names
        else:
            value, names = x, (None,) * x.dim()
        if axis_name in names:
            log.warning("Param %s with names %s already sharded on axis %s", tuple(value.shape), names, axis_name)
            return x
        if value.numel() <= min_weight_size:
            log.info("Param %s too small to shard (<= %d)", tuple(value.shape), min_weight_size)
            return x
        i = choose_shard_dim(tuple(value.shape), names, axis_size)
        if i is None:
            log.warning("could not shard %s on axis %s: no suitable dim found", tuple(value.shape), axis_name)
            return x
        split = value.shape[i] // axis_size
        sl = value.narrow(i, axis_idx * split, split).contiguous()
        return Partitioned(value=sl, names=names[:i] + (axis_name,) + names[i + 1:])

    return tree_map(_split, params, is_leaf=is_partitioned)


class _GatherMeanGrad(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, axis, axis_name, mesh):
        ctx.meta = (axis, axis_name, mesh)
        return C.all_gather(x, axis_name, axis=axis, tiled=True, mesh=mesh)

    @staticmethod
    def backward(ctx, g):
        axis, axis_name, mesh = ctx.meta
        n = mesh.axis_size(axis_name)
        return C.psum_scatter(g.contiguous(), axis_name, scatter_dimension=axis, tiled=True, mesh=mesh) / n, None, None, None


def gather_arr_mean_grads(x: torch.Tensor, axis: int, axis_name: str, mesh: Optional[Mesh] = None) -> torch.Tensor:
    """Forward: tiled all-gather along ``axis``. Backward: reduce-scatter of the gradient
    divided by the axis size, i.e. the shard receives the mean over devices of the gradient of
    its slice (param_sharding.py:129-142)."""
    m = mesh or current_mesh()
    return _GatherMeanGrad.apply(x, axis, axis_name, m)


@named_scope("gather_params")
def gather_params(params: Pytree, axis_name: str, mesh: Optional[Mesh] = None) -> Pytree:
    """Gathers every leaf sharded on ``axis_name``; keeps the box only if other axis names
    remain, so it composes with further mesh axes (param_sharding.py:146-175)."""
    m = mesh or current_mesh()

    def _gather(p):
        if is_partitioned(p) and axis_name in p.names:
            shard_axis = p.names.index(axis_name)
            value = gather_arr_mean_grads(p.value, axis=shard_axis, axis_name=axis_name, mesh=m)
            names = p.names[:shard_axis] + (None,) + p.names[shard_axis + 1:]
            if any(n is not None for n in names):
                return Partitioned(value, names)
            return value
        return p

    return tree_map(_gather, params, is_leaf=is_partitioned)


class _ShardedModule(Module):
    """``nn.map_variables(target, trans_in_fn=gather_params, trans_out_fn=shard_params)``:
    parameters are stored sharded, gathered on every read, re-sharded on write
    (param_sharding.py:179-191)."""

    def __init__(self, inner: Module, axis_name: str, min_weight_size: int):
        self.inner, self.axis_name, self.min_weight_size = inner, axis_name, min_weight_size
        self.name = inner.name

    def init_params(self, key, *args, **kw):
        full = self.inner.init_params(key, *args, **kw)
        return shard_params(full, self.axis_name, self.min_weight_size)

    def forward(self, params, *args, **kw):
        return self.inner.forward(gather_params(params, self.axis_name), *args, **kw)


def shard_module_params(target, axis_name: str, min_weight_size: int = 2 ** 18):
    """Wraps a module instance, or a module class/factory (returns a factory), so that its
    parameters live sharded over ``axis_name`` (param_sharding.py:179-191)."""
    if isinstance(target, Module):
        return _ShardedModule(target, axis_name, min_weight_size)

    def factory(*a, **kw):
        return _ShardedModule(target(*a, **kw), axis_name, min_weight_size)

    return factory


def init_model(rng: _rng.Key, x: torch.Tensor, model, learning_rate: float = 1e-4, optimizer=None) -> TrainState:
    """param_sharding.py:227-241: every rank initialises the full parameter from the same rng and
    keeps its slice; AdamW state is created on the sharded leaves and inherits their boxes."""
    init_rng, rng = _rng.split(rng, 2)
    variables = model.init({"params": init_rng}, x, train=False)
    params = variables.pop("params")
    tx = optimizer if optimizer is not None else adamw(learning_rate)
    # the stored key is always in rebased form (salt folded into the 64-bit state), the same
    # form every train step leaves it in
    return TrainState.create(apply_fn=model.apply, params=params, tx=tx, rng=_rng.rebase(rng))


def sync_gradients(grads: Pytree, axis_names: Sequence[str], mesh: Optional[Mesh] = None) -> Pytree:
    """Averages gradients over the axes a parameter is *replicated* on: sharded leaves skip the
    axes in their ``names``; plain leaves are averaged over all ``axis_names``
    (param_sharding.py:293-322). N-D aware."""
    m = mesh or current_mesh()
    axis_names = tuple(axis_names)

    def _sync(g):
        if is_partitioned(g):
            repl = [a for a in axis_names if a not in g.names]
            if not repl:
                return g
            return g.replace(value=C.pmean(g.value, axis_name=repl, mesh=m))
        return C.pmean(g, axis_name=axis_names, mesh=m)

    return tree_map(_sync, grads, is_leaf=is_partitioned)


def train_step_fsdp(state: TrainState, metrics: Optional[Metrics], batch: Batch, *, num_minibatches: int = 4,
                    data_axis_name: str = "data", loss_fn: Optional[Callable] = None,
                    sync_axes: Optional[Sequence[str]] = None) -> Tuple[TrainState, Metrics]:
    """param_sharding.py:343-367."""
    loss_fn = loss_fn or make_loss_fn(data_axis_name)
    rng, step_rng = _rng.split(state.rng, 2)
    grads, step_metrics = accum_grads(state, batch, step_rng, num_minibatches, loss_fn=loss_fn)
    with named_scope("sync_grad"):
        grads = sync_gradients(grads, tuple(sync_axes) if sync_axes else (data_axis_name,))
    new_state = state.apply_gradients(grads=grads, rng=_rng.rebase(rng))
    with named_scope("synch_metrics"):
        axes = tuple(sync_axes) if sync_axes else data_axis_name
        step_metrics = tree_map(lambda x: C.psum(x, axis_name=axes), step_metrics)
    if metrics is None:
        metrics = step_metrics
    else:
        metrics = tree_map(lambda a, b: a + b, metrics, step_metrics)
    return new_state, metrics


def make_train_step_fsdp(mesh: Optional[Mesh] = None, **kw) -> Callable:
    """The analogue of ``jax.jit(shard_map(train_step_fsdp, in_specs=(state_specs, P(), P(data)),
    ...), donate_argnames=...)`` (param_sharding.py:370-379)."""
    mesh = mesh or current_mesh()

    def step(state: TrainState, metrics: Optional[Metrics], batch: Batch):
        if state.engine is not None:
            return state.engine.train_step(state, metrics, batch)
        with mesh:
            return train_step_fsdp(state, metrics, batch, **kw)

    return step


def unshard_params(params: Pytree, mesh: Optional[Mesh] = None) -> Pytree:
    """Full (global) arrays of a sharded tree, gathered over every named axis — the view a
    checkpoint stores. No gradient."""
    m = mesh or current_mesh()

    def _full(p):
        if not is_partitioned(p):
            return p
        v = p.value
        for d, name in enumerate(p.names):
            if name is not None:
                v = C.all_gather(v.detach(), name, axis=d, tiled=True, mesh=m)
        return v

    return tree_map(_full, params, is_leaf=is_partitioned)

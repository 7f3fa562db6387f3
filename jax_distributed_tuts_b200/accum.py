"""Gradient accumulation over minibatches (reference: util.py:41-167).

``accum_grads_loop`` is the reference's live path: a Python loop over ``n_minbatch`` equal
slices, per-minibatch RNG from ``split(key, n)``, gradients *averaged* (``/ n_minbatch``,
util.py:77) and metrics *summed* (util.py:75).

``accum_grads_scan`` is the reference's ``lax.scan`` variant (one compiled body instead of n
unrolled copies, util.py:81-137): zero-initialised carries with the shapes of one minibatch
step (util.py:123-129) and the same body applied to ``narrow``-ed windows of the batch
(the ``dynamic_slice_in_dim`` of util.py:97-105). On CPU the body simply runs n times. On a GPU
the body (forward + backward of one minibatch + accumulation into the carry) is captured ONCE as
a CUDA graph over static buffers and replayed n times per step and on every later step
(``ScanBody``): the "compile one body instead of n" of the reference without a tracing compiler.
Because a replayed graph cannot take a different host-side RNG salt per minibatch, the captured
body derives minibatch i's stream on the device (``rng.rebase(keys[i])`` written into a fixed
key buffer): dropout masks are a different, equally valid per-minibatch stream than the loop's;
with dropout off the two variants agree bit for bit. The fused engines go further: their whole
step is one graph and the minibatch loop is folded into the GEMMs
(``engine.MLPEngine(fuse_minibatches=True)``).
"""
from __future__ import annotations

from typing import Callable, Tuple

import torch

from . import rng as _rng
from .autodiff import value_and_grad
from .state import Batch, Metrics, TrainState
from .tree import Pytree, is_partitioned, tree_leaves, tree_map


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


class ScanBody:
    """One minibatch step (``value_and_grad(loss_fn)`` + accumulation into the carry) captured as a
    CUDA graph. Static buffers: a copy of the parameters, one minibatch of inputs / labels, the RNG
    key state, the gradient and metric carries. ``run`` = zero the carries, then per minibatch copy
    the window into the static batch, write its key, replay."""

    def __init__(self, state: TrainState, example: Batch, loss_fn: Callable):
        dev = example.inputs.device
        clone = lambda p: p.replace(value=p.value.detach().clone()) if is_partitioned(p) else p.detach().clone()
        self.params = tree_map(clone, state.params, is_leaf=is_partitioned)
        self.inputs, self.labels = example.inputs.clone(), example.labels.clone()
        self.key_buf = torch.zeros(1, dtype=torch.int64, device=dev)
        self.apply_fn, self.loss_fn = state.apply_fn, loss_fn
        self.grad_fn = value_and_grad(loss_fn, has_aux=True)
        zeros = lambda x: (x.replace(value=torch.zeros_like(x.value)) if is_partitioned(x) else torch.zeros_like(x))
        self.grads = tree_map(zeros, self.params, is_leaf=is_partitioned)
        self.metric_tensors = None      # flat list of accumulated tensor-valued metric entries
        self.metric_template = None
        # warm-up on a side stream (lazy initialisation must not happen inside the capture), then capture
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self._body(first=True)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._body(first=False)
        self.replays = 0

    def _body(self, first: bool) -> None:
        key = _rng.Key(self.key_buf, 0)
        (_, step_metric), step_grad = self.grad_fn(self.params, self.apply_fn, Batch(self.inputs, self.labels), key)
        flat = []

        def collect(x):
            if isinstance(x, torch.Tensor):
                flat.append(x)
            return x
        tree_map(collect, step_metric)
        if first:
            self.metric_template = step_metric
            self.metric_tensors = [torch.zeros_like(t) for t in flat]
            return
        tree_map(lambda g, sg: (g.value if is_partitioned(g) else g).add_(sg.value if is_partitioned(sg) else sg),
                 self.grads, step_grad, is_leaf=is_partitioned)
        for acc, t in zip(self.metric_tensors, flat):
            acc.add_(t)

    def run(self, batch: Batch, state: TrainState, keys, n_minbatch: int) -> Tuple[Pytree, Metrics]:
        mbs = self.inputs.shape[0]
        tree_map(lambda dst, src: (dst.value if is_partitioned(dst) else dst).copy_(src.value if is_partitioned(src) else src),
                 self.params, state.params, is_leaf=is_partitioned)
        tree_map(lambda g: (g.value if is_partitioned(g) else g).zero_(), self.grads, is_leaf=is_partitioned)
        for t in self.metric_tensors:
            t.zero_()
        for i in range(n_minbatch):
            self.inputs.copy_(batch.inputs.narrow(0, i * mbs, mbs))
            self.labels.copy_(batch.labels.narrow(0, i * mbs, mbs))
            self.key_buf.copy_(_rng.rebase(keys[i]).base)
            self.graph.replay()
            self.replays += 1
        grads = tree_map(lambda g: g.replace(value=g.value / n_minbatch) if is_partitioned(g) else g / n_minbatch,
                         self.grads, is_leaf=is_partitioned)
        it = iter(self.metric_tensors)
        # tensor entries come from the accumulators, host numbers (row counts) are multiplied out
        metrics = tree_map(lambda x: next(it).clone() if isinstance(x, torch.Tensor) else x * n_minbatch, self.metric_template)
        return grads, metrics


_SCAN_CACHE: dict = {}
_SCAN_CACHE_MAX = 8


def accum_grads_scan(batch: Batch, state: TrainState, key: _rng.Key, n_minbatch: int,
                     loss_fn: Callable) -> Tuple[Pytree, Metrics]:
    bs = batch.inputs.shape[0]
    min_batch_size = bs // n_minbatch
    keys = _rng.split(key, n_minbatch)
    if batch.inputs.is_cuda and key.base.is_cuda:
        leaves = [p.value if is_partitioned(p) else p for p in tree_leaves(state.params, is_leaf=is_partitioned)]
        sig = (id(loss_fn), id(state.apply_fn), min_batch_size, tuple(batch.inputs.shape[1:]), batch.inputs.dtype,
               batch.labels.dtype, tuple((tuple(t.shape), t.dtype) for t in leaves), str(batch.inputs.device))
        body = _SCAN_CACHE.get(sig)
        if body is None:
            # the entry keeps loss_fn / apply_fn alive, so their ids in `sig` cannot be recycled while it exists
            body = ScanBody(state, batch.slice(0, min_batch_size), loss_fn)
            if len(_SCAN_CACHE) >= _SCAN_CACHE_MAX:
                _SCAN_CACHE.pop(next(iter(_SCAN_CACHE)))       # oldest captured body (and its static buffers)
            _SCAN_CACHE[sig] = body
        return body.run(batch, state, keys, n_minbatch)
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

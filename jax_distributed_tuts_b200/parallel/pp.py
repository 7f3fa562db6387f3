"""Pipeline parallelism: GPipe with microbatches shifted stage to stage.

The reference's ``pipeline_parallel.py`` is a stub (imports and three type aliases,
pipeline_parallel.py:1-39); its imports say what the train step was meant to reuse — ``Batch``,
``TrainState``, gradient accumulation, ``fold_rng_over_axis`` (per-stage RNG, :37) and
``sync_gradients`` (gradient mean over the non-sharded mesh axes, :38) — and BASELINE.json fixes
the algorithm: layers split into stages, the batch split into microbatches, activations moved
stage -> stage+1 (``ppermute``) for M + S - 1 ticks, loss on the last stage, backward in reverse,
then gradient sync + optimizer per stage.

Here every rank of the ``pipe`` mesh axis owns a contiguous range of layers. Activations and
activation gradients travel as NCCL point-to-point messages (``batch_isend_irecv``) posted ahead
of the compute that needs them, so the transfer of microbatch m+1 overlaps the compute of
microbatch m; compute is our kernels behind autograd (generic tier; the product path with per-stage fused engines and CUDA
graphs is ``engine/pp.py``). A ``data`` axis, if the mesh has one, averages gradients with ``sync_gradients``
semantics.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Tuple

import torch

from .. import rng as _rng
from ..comm import collectives as C
from ..mesh import Mesh, current_mesh
from ..models.gpt2 import GPT2
from ..models.mlp import Module
from ..ops import functional as Fn
from ..optim import adamw
from ..state import Batch, Metrics, TrainState
from ..tree import tree_leaves, tree_map
from ..utils.profiling import named_scope
from .dp import fold_rng_over_axis
from .fsdp import sync_gradients


def split_layers(n_layer: int, n_stages: int) -> List[range]:
    """Contiguous, as-even-as-possible layer ranges (earlier stages take the remainder)."""
    base, rem = divmod(n_layer, n_stages)
    out, start = [], 0
    for s in range(n_stages):
        n = base + (1 if s < rem else 0)
        out.append(range(start, start + n))
        start += n
    return out


class GPT2Stage(Module):
    """The slice of a GPT-2 that one pipeline stage owns. Stage 0 additionally owns the
    embeddings, the last stage the final norm and a copy of ``wte`` for the tied LM head (the two
    copies receive the sum of both gradients each step, see ``train_step_pp``)."""

    def __init__(self, model: GPT2, stage: int, n_stages: int):
        self.model, self.stage, self.n_stages = model, stage, n_stages
        self.layers = split_layers(model.config.n_layer, n_stages)[stage]
        self.first, self.last = stage == 0, stage == n_stages - 1

    def init_params(self, key: _rng.Key, tokens: torch.Tensor, **kw) -> Dict:
        full = self.model.init_params(key, tokens)      # same rng on every stage -> same values; keep our slice
        p: Dict = {f"h_{i}": full[f"h_{i}"] for i in self.layers}
        if self.first:
            p["wte"], p["wpe"] = full["wte"], full["wpe"]
        if self.last:
            p["ln_f"] = full["ln_f"]
            if not self.first:
                p["wte"] = full["wte"].clone()
        return p

    def forward(self, params: Dict, x: torch.Tensor, train: bool = False, rngs=None, **kw) -> torch.Tensor:
        if self.first:
            x = self.model.embed(params, x)
        for i in self.layers:
            x = self.model.block(params[f"h_{i}"], x)
        if self.last:
            x = self.model.head(params, x).float()
        return x

    def act_shape(self, batch: int, seq: int) -> Tuple[int, ...]:
        return (batch, seq, self.model.config.d_model)


def init_pp(rng: _rng.Key, tokens: torch.Tensor, stage_model: Module, learning_rate: float = 3e-4,
            optimizer=None) -> TrainState:
    """Per-stage TrainState; all stages use the same rng so a pipelined model starts from exactly
    the parameters of the unsplit one (the reference's init convention, data_paral.py:128-147)."""
    init_rng, rng = _rng.split(rng, 2)
    params = stage_model.init({"params": init_rng}, tokens, train=False).pop("params")
    tx = optimizer if optimizer is not None else adamw(learning_rate)
    return TrainState.create(apply_fn=stage_model.apply, params=params, tx=tx, rng=_rng.rebase(rng))


def _xent_metrics(logits: torch.Tensor, labels: torch.Tensor):
    v = logits.shape[-1]
    loss, correct = Fn.softmax_cross_entropy_with_integer_labels(logits.reshape(-1, v), labels.reshape(-1),
                                                                 return_correct=True)
    n = labels.numel()
    return loss.mean(), {"loss": (loss.sum().detach(), n), "accuracy": (correct.detach(), n)}


def train_step_pp(state: TrainState, metrics: Optional[Metrics], batch: Batch, *, stage_model: GPT2Stage,
                  num_microbatches: int = 4, pipe_axis: str = "pipe", data_axis: Optional[str] = None,
                  mesh: Optional[Mesh] = None) -> Tuple[TrainState, Metrics]:
    """One GPipe step on this stage. ``batch`` is the (data-axis-local) batch: stage 0 reads
    ``batch.inputs``, the last stage ``batch.labels``; every stage sees the same batch object
    (only shapes are used elsewhere), as under the reference's ``shard_map``."""
    m = mesh or current_mesh()
    S, s = m.axis_size(pipe_axis), m.axis_index(pipe_axis)
    first, last = s == 0, s == S - 1
    dev = tree_leaves(state.params)[0].device
    rng, step_rng = _rng.split(state.rng, 2)
    step_rng = fold_rng_over_axis(step_rng, pipe_axis, m)           # distinct stream per stage
    B, T = batch.inputs.shape[0], batch.inputs.shape[1]
    assert B % num_microbatches == 0
    mb = B // num_microbatches
    act_shape = stage_model.act_shape(mb, T)

    # parameters as autograd leaves for this step
    leaves = tree_map(lambda p: p.detach().requires_grad_(True), state.params)
    flat = tree_leaves(leaves)

    def post_recv(src: int):
        buf = torch.empty(act_shape, dtype=torch.bfloat16, device=dev)
        return buf, C.send_recv(None, None, buf, src, pipe_axis, m)

    def wait(reqs):
        for r in reqs:
            r.wait()

    # ---------------- forward: all microbatches ----------------
    saved: List[Tuple[Optional[torch.Tensor], torch.Tensor]] = []
    step_metrics = None
    pending_sends: List = []
    nxt = post_recv(s - 1) if not first else None
    with torch.enable_grad():
        for k in range(num_microbatches):
            sl = slice(k * mb, (k + 1) * mb)
            if first:
                x_in, inp = None, batch.inputs[sl]
            else:
                buf, reqs = nxt
                wait(reqs)
                nxt = post_recv(s - 1) if k + 1 < num_microbatches else None    # prefetch the next microbatch
                x_in = buf.requires_grad_(True)
                inp = x_in
            with named_scope("pp_stage_forward"):
                out = state.apply_fn({"params": leaves}, inp, train=True, rngs={"dropout": _rng.fold_in(step_rng, k)})
            if last:
                loss, mm = _xent_metrics(out, batch.labels[sl])
                step_metrics = mm if step_metrics is None else tree_map(lambda a, b: a + b, step_metrics, mm)
                saved.append((x_in, loss / num_microbatches))
            else:
                send = out.detach().to(torch.bfloat16).contiguous()
                pending_sends.append((send, C.send_recv(send, s + 1, None, None, pipe_axis, m)))
                saved.append((x_in, out))
    # (the forward sends stay in flight: they are only waited for at the end of the step, together with the
    #  backward ones, so a slow receiver never stalls this stage between its two phases)

    # ---------------- backward: microbatches in reverse ----------------
    nxt = post_recv(s + 1) if not last else None
    for k in range(num_microbatches - 1, -1, -1):
        x_in, out = saved[k]
        if last:
            grad_out = None
        else:
            buf, reqs = nxt
            wait(reqs)
            nxt = post_recv(s + 1) if k > 0 else None
            grad_out = buf.to(out.dtype)
        with named_scope("pp_stage_backward"):
            inputs = flat + ([x_in] if x_in is not None else [])
            grads = torch.autograd.grad(out, inputs, grad_outputs=grad_out, allow_unused=True)
        for p, g in zip(flat, grads[: len(flat)]):
            if g is not None:
                p.grad = g if p.grad is None else p.grad + g
        if x_in is not None:
            gx = grads[-1].to(torch.bfloat16).contiguous()
            pending_sends.append((gx, C.send_recv(gx, s - 1, None, None, pipe_axis, m)))
        saved[k] = (None, None)
    for _, reqs in pending_sends:
        wait(reqs)

    grads_tree = tree_map(lambda p: p.grad if p.grad is not None else torch.zeros_like(p), leaves)

    # tied embedding: first and last stage exchange their wte gradients and both keep the sum
    if S > 1 and (first or last) and "wte" in grads_tree:
        mine = grads_tree["wte"].contiguous()
        other = torch.empty_like(mine)
        peer = S - 1 if first else 0
        wait(C.send_recv(mine, peer, other, peer, pipe_axis, m))
        grads_tree = dict(grads_tree, wte=mine + other)

    if data_axis is not None:
        with named_scope("sync_grad"):
            grads_tree = sync_gradients(grads_tree, (data_axis,), m)
    new_state = state.apply_gradients(grads=grads_tree, rng=_rng.rebase(rng))

    with named_scope("sync_metrics"):
        if step_metrics is None:   # not the last stage: contribute zeros, counts included
            z = torch.zeros((), device=dev)
            step_metrics = {"loss": (z, 0), "accuracy": (z.clone(), 0)}
        axes = (pipe_axis,) + ((data_axis,) if data_axis else ())
        step_metrics = tree_map(lambda x: C.psum(x, axis_name=axes, mesh=m), step_metrics)
    if metrics is None:
        metrics = step_metrics
    else:
        metrics = tree_map(lambda a, b: a + b, metrics, step_metrics)
    return new_state, metrics


def make_train_step_pp(mesh: Optional[Mesh] = None, **kw) -> Callable:
    mesh = mesh or current_mesh()

    def step(state: TrainState, metrics: Optional[Metrics], batch: Batch):
        with mesh:
            return train_step_pp(state, metrics, batch, mesh=mesh, **kw)

    return step

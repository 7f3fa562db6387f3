"""Collectives addressed by mesh-axis name (the ``jax.lax`` surface the reference uses).

    axis_index / axis_size   data_paral.py:33, param_sharding.py:63,65,130
    psum / pmean             data_paral.py:212,223, param_sharding.py:313,318,358
    all_gather(tiled)        param_sharding.py:140
    psum_scatter(tiled)      param_sharding.py:136
    ppermute                 intended by pipeline_parallel.py (absent there)

Three transports sit behind the same calls:

    "fused"  one sm_100a kernel per collective over NVSwitch peer / multicast memory
             (``comm/fused.py`` + ``csrc/comm.cu``: flag exchange, ``multimem.ld_reduce``,
             ``multimem.st``), on a symmetric workspace owned by the mesh axis. GPUs only.
    "nccl"   torch.distributed on the axis' NCCL communicator (the baseline transport).
    "gloo"   torch.distributed on CPU processes (the fake-cluster mode, util.py:31-38).

``set_backend("auto")`` (default) uses NCCL on GPUs and gloo on CPU; ``set_backend("fused")`` or
``B200DIST_COLLECTIVES=fused`` switches psum / pmean / psum_scatter / all_gather on GPUs to the
fused kernels (anything they do not handle, and ppermute / send_recv, stay on NCCL: pipeline
traffic is point-to-point by design). The fused *optimizer* paths (gradient all-reduce + AdamW,
reduce-scatter + sharded AdamW, all-gather inside the GEMM) are in ``engine/``.
"""
from __future__ import annotations

import os
from typing import Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from ..mesh import AxisNames, Mesh, current_mesh
from ..tree import Pytree, tree_map


_BACKEND = os.environ.get("B200DIST_COLLECTIVES", "auto")
fused_calls = 0      # number of collectives served by the fused kernels (tests / diagnostics)


def set_backend(name: str) -> None:
    """"auto" | "fused" | "nccl" | "gloo" (the last two both mean: torch.distributed on the axis group)."""
    global _BACKEND
    assert name in ("auto", "fused", "nccl", "gloo"), name
    _BACKEND = name


def get_backend() -> str:
    return _BACKEND


def _use_fused(m: Mesh) -> bool:
    if _BACKEND != "fused":
        return False
    from . import fused
    return fused.available(m)


def _mesh(mesh: Optional[Mesh]) -> Mesh:
    return mesh if mesh is not None else current_mesh()


def axis_index(axis_name: AxisNames, mesh: Optional[Mesh] = None) -> int:
    return _mesh(mesh).axis_index(axis_name)


def axis_size(axis_name: AxisNames, mesh: Optional[Mesh] = None) -> int:
    return _mesh(mesh).axis_size(axis_name)


def _as_tensor(x, device) -> Tuple[torch.Tensor, bool]:
    if isinstance(x, torch.Tensor):
        return x, True
    return torch.tensor(x, device=device), False


def _psum_leaf(x, axis_name, mesh: Mesh, mean: bool):
    n = mesh.axis_size(axis_name)
    if n == 1:
        return x
    t, was_tensor = _as_tensor(x, mesh.device)
    if _use_fused(mesh):
        from . import fused
        got = fused.psum(t, axis_name, mesh, mean)
        if got is not None:
            global fused_calls
            fused_calls += 1
            return got if was_tensor else got.item()
    out = t.clone() if was_tensor else t
    if out.dtype in (torch.bfloat16, torch.float16) and not out.is_cuda:
        tmp = out.float()
        dist.all_reduce(tmp, op=dist.ReduceOp.SUM, group=mesh.group(axis_name))
        out = tmp.to(out.dtype)
    else:
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=mesh.group(axis_name))
    if mean:
        out = out / n
    return out if was_tensor else out.item()


def psum(x: Pytree, axis_name: AxisNames, mesh: Optional[Mesh] = None) -> Pytree:
    m = _mesh(mesh)
    return tree_map(lambda v: _psum_leaf(v, axis_name, m, False), x)


def pmean(x: Pytree, axis_name: AxisNames, mesh: Optional[Mesh] = None) -> Pytree:
    m = _mesh(mesh)
    return tree_map(lambda v: _psum_leaf(v, axis_name, m, True), x)


def all_gather(x: torch.Tensor, axis_name: AxisNames, *, axis: int = 0, tiled: bool = True,
               mesh: Optional[Mesh] = None) -> torch.Tensor:
    """Concatenates the per-rank ``x`` along ``axis`` (``tiled=True``) or stacks on a new axis."""
    m = _mesh(mesh)
    n = m.axis_size(axis_name)
    if n == 1:
        return x if tiled else x.unsqueeze(axis)
    xc = x.contiguous()
    out = None
    if _use_fused(m):
        from . import fused
        out = fused.all_gather(xc, axis_name, m)
        if out is not None:
            global fused_calls
            fused_calls += 1
    if out is None:
        out = torch.empty((n,) + tuple(xc.shape), dtype=xc.dtype, device=xc.device)
        dist.all_gather_into_tensor(out.view(-1), xc.view(-1), group=m.group(axis_name))
    if not tiled:
        return out.movedim(0, axis)
    # [n, ..., d_axis, ...] -> [..., n*d_axis, ...]
    out = out.movedim(0, axis)
    shape = list(x.shape)
    shape[axis] *= n
    return out.reshape(shape)


def psum_scatter(x: torch.Tensor, axis_name: AxisNames, *, scatter_dimension: int = 0, tiled: bool = True,
                 mesh: Optional[Mesh] = None) -> torch.Tensor:
    """Sum over the axis, each rank keeping its slice of ``scatter_dimension``."""
    m = _mesh(mesh)
    n = m.axis_size(axis_name)
    if n == 1:
        return x
    assert tiled, "only tiled=True is used by the strategies"
    d = scatter_dimension
    assert x.shape[d] % n == 0, f"dim {d} of {tuple(x.shape)} not divisible by axis size {n}"
    # bring the scattered dim to the front in n chunks
    chunks = x.movedim(d, 0).reshape((n, x.shape[d] // n) + tuple(s for i, s in enumerate(x.shape) if i != d))
    chunks = chunks.contiguous()
    if _use_fused(m):
        from . import fused
        got = fused.psum_scatter(chunks, axis_name, m)
        if got is not None:
            global fused_calls
            fused_calls += 1
            return got.movedim(0, d)
    out = torch.empty_like(chunks[0])
    if chunks.is_cuda:
        dist.reduce_scatter_tensor(out.view(-1), chunks.view(-1), op=dist.ReduceOp.SUM, group=m.group(axis_name))
    else:  # gloo has no reduce_scatter: all-reduce then slice
        tmp = chunks.float() if chunks.dtype in (torch.bfloat16, torch.float16) else chunks.clone()
        dist.all_reduce(tmp, op=dist.ReduceOp.SUM, group=m.group(axis_name))
        out = tmp[m.axis_index(axis_name)].to(x.dtype)
    return out.movedim(0, d)


def send_recv(send: Optional[torch.Tensor], dst_index: Optional[int], recv: Optional[torch.Tensor],
              src_index: Optional[int], axis_name: str, mesh: Optional[Mesh] = None) -> list:
    """Posts a paired isend/irecv along one mesh axis (pipeline stage-to-stage traffic).
    Returns the request handles; callers ``wait()`` on them when the data is needed."""
    m = _mesh(mesh)
    ranks = m.group_ranks(axis_name)
    ops = []
    if send is not None and dst_index is not None:
        ops.append(dist.P2POp(dist.isend, send, ranks[dst_index]))
    if recv is not None and src_index is not None:
        ops.append(dist.P2POp(dist.irecv, recv, ranks[src_index]))
    if not ops:
        return []
    return dist.batch_isend_irecv(ops)


def ppermute(x: torch.Tensor, axis_name: str, perm: Sequence[Tuple[int, int]],
             mesh: Optional[Mesh] = None) -> torch.Tensor:
    """``jax.lax.ppermute``: rank ``src`` sends ``x`` to rank ``dst`` for every (src, dst) in
    ``perm``; ranks that receive nothing get zeros."""
    m = _mesh(mesh)
    n = m.axis_size(axis_name)
    if n == 1:
        return x if any(s == 0 and d == 0 for s, d in perm) else torch.zeros_like(x)
    me = m.axis_index(axis_name)
    dst = next((d for s, d in perm if s == me), None)
    src = next((s for s, d in perm if d == me), None)
    out = torch.zeros_like(x)
    if dst == me and src == me:
        return x.clone()
    reqs = send_recv(x.contiguous() if dst is not None else None, dst, out if src is not None else None, src,
                     axis_name, m)
    for r in reqs:
        r.wait()
    return out

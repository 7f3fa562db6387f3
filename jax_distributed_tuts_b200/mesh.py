"""Named-axis device mesh over one-process-per-GPU ranks.

Replaces ``jax.sharding.Mesh(np.array(jax.devices()), ("data",))`` + ``shard_map``
(reference: data_paral.py:150-163, param_sharding.py:249-274). Under ``shard_map`` the body
runs once per device and names collectives by mesh axis; here each *process* is one mesh
coordinate and the mesh owns one communicator per axis subset (NCCL on GPUs, gloo on CPU),
so strategy code is written once against axis names.
"""
from __future__ import annotations

import itertools
from typing import Dict, Optional, Sequence, Tuple, Union

import torch
import torch.distributed as dist

AxisNames = Union[str, Sequence[str]]
_CURRENT: list["Mesh"] = []


class Mesh:
    def __init__(self, shape: Sequence[int], axis_names: Sequence[str], *, rank: Optional[int] = None,
                 world_size: Optional[int] = None, device: Optional[torch.device] = None):
        self.shape = tuple(int(s) for s in shape)
        self.axis_names = tuple(axis_names)
        assert len(self.shape) == len(self.axis_names) and len(set(self.axis_names)) == len(self.axis_names)
        n = 1
        for s in self.shape:
            n *= s
        self.size = n
        if dist.is_available() and dist.is_initialized():
            self.rank = dist.get_rank() if rank is None else rank
            ws = dist.get_world_size() if world_size is None else world_size
        else:
            self.rank = 0 if rank is None else rank
            ws = 1 if world_size is None else world_size
        if ws != n:
            raise ValueError(f"mesh {dict(zip(self.axis_names, self.shape))} needs {n} ranks, world has {ws}")
        self.coords = self._coords_of(self.rank)
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if (
                torch.cuda.is_available() and dist.is_initialized() and dist.get_backend() == "nccl"
            ) else (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() and n == 1
                    else torch.device("cpu"))
        self.device = device
        self._groups: Dict[Tuple[str, ...], Optional[dist.ProcessGroup]] = {}
        self._group_ranks: Dict[Tuple[str, ...], Tuple[int, ...]] = {}
        self._build_groups()

    # -- coordinates -------------------------------------------------------------------------
    def _coords_of(self, rank: int) -> Tuple[int, ...]:
        c = []
        for s in reversed(self.shape):
            c.append(rank % s)
            rank //= s
        return tuple(reversed(c))

    def _rank_of(self, coords: Sequence[int]) -> int:
        r = 0
        for c, s in zip(coords, self.shape):
            r = r * s + c
        return r

    def _norm(self, axes: AxisNames) -> Tuple[str, ...]:
        if isinstance(axes, str):
            axes = (axes,)
        axes = tuple(axes)
        for a in axes:
            if a not in self.axis_names:
                raise KeyError(f"unknown mesh axis {a!r}; mesh axes are {self.axis_names}")
        return tuple(a for a in self.axis_names if a in axes)  # canonical order

    def _build_groups(self) -> None:
        """One communicator per non-empty subset of axes (every rank creates every group, in the
        same order, as torch.distributed requires)."""
        nd = len(self.axis_names)
        for r in range(1, nd + 1):
            for subset in itertools.combinations(range(nd), r):
                names = tuple(self.axis_names[i] for i in subset)
                others = [i for i in range(nd) if i not in subset]
                mine_ranks: Tuple[int, ...] = ()
                mine_group = None
                for fixed in itertools.product(*(range(self.shape[i]) for i in others)):
                    ranks = []
                    for var in itertools.product(*(range(self.shape[i]) for i in subset)):
                        coords = [0] * nd
                        for i, v in zip(others, fixed):
                            coords[i] = v
                        for i, v in zip(subset, var):
                            coords[i] = v
                        ranks.append(self._rank_of(coords))
                    g = None
                    if self.size > 1 and len(ranks) > 1:
                        g = dist.group.WORLD if len(ranks) == self.size else dist.new_group(ranks)
                    if self.rank in ranks:
                        mine_ranks, mine_group = tuple(ranks), g
                self._groups[names] = mine_group
                self._group_ranks[names] = mine_ranks

    # -- queries (the analogues of lax.axis_index / psum(1, axis)) ----------------------------
    def axis_index(self, axes: AxisNames) -> int:
        names = self._norm(axes)
        idx = 0
        for a in names:
            i = self.axis_names.index(a)
            idx = idx * self.shape[i] + self.coords[i]
        return idx

    def axis_size(self, axes: AxisNames) -> int:
        n = 1
        for a in self._norm(axes):
            n *= self.shape[self.axis_names.index(a)]
        return n

    def group(self, axes: AxisNames) -> Optional[dist.ProcessGroup]:
        return self._groups[self._norm(axes)]

    def group_ranks(self, axes: AxisNames) -> Tuple[int, ...]:
        return self._group_ranks[self._norm(axes)]

    # -- context --------------------------------------------------------------------------------
    def __enter__(self) -> "Mesh":
        _CURRENT.append(self)
        return self

    def __exit__(self, *exc) -> None:
        _CURRENT.pop()

    def __repr__(self) -> str:
        return f"Mesh({dict(zip(self.axis_names, self.shape))}, rank={self.rank}, coords={self.coords})"


def set_mesh(mesh: Optional[Mesh]) -> None:
    """Installs a process-wide default mesh (what the module-level ``mesh`` is in the reference)."""
    _CURRENT.clear()
    if mesh is not None:
        _CURRENT.append(mesh)


def current_mesh() -> Mesh:
    if not _CURRENT:
        raise RuntimeError("no active mesh: use `with mesh:` or set_mesh(mesh)")
    return _CURRENT[-1]


def make_mesh(axis_names: Sequence[str] = ("data",), shape: Optional[Sequence[int]] = None) -> Mesh:
    """Mesh over all ranks of the job; 1-D by default like the reference's ``("data",)`` mesh."""
    ws = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    if shape is None:
        shape = (ws,) + (1,) * (len(axis_names) - 1)
    return Mesh(shape, axis_names)

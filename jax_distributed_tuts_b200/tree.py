"""Minimal pytree utilities and the ``Partitioned`` box.

Pytrees are nested dicts / lists / tuples / dataclass instances registered with
:func:`register_dataclass`. ``Partitioned(value, names)`` plays the role of
``flax.linen.Partitioned`` (reference: param_sharding.py:72-113,146-175): it records, per
tensor dimension, the mesh axis the *stored local slice* is sharded over.
"""
from __future__ import annotations

import dataclasses
from typing import Any, Callable, Iterable, Optional, Tuple

import torch

Pytree = Any
_DATACLASSES: dict[type, Tuple[Tuple[str, ...], Tuple[str, ...]]] = {}


def register_dataclass(cls=None, *, static: Iterable[str] = ()):
    """Registers a dataclass as a pytree node; ``static`` fields are carried, not mapped."""
    def wrap(c):
        static_f = tuple(static)
        data_f = tuple(f.name for f in dataclasses.fields(c) if f.name not in static_f)
        _DATACLASSES[c] = (data_f, static_f)
        return c
    return wrap(cls) if cls is not None else wrap


@dataclasses.dataclass(frozen=True)
class Partitioned:
    """A parameter (local slice) plus the mesh-axis name of each of its dimensions."""
    value: Any
    names: Tuple[Optional[str], ...]

    def replace(self, **kw) -> "Partitioned":
        return dataclasses.replace(self, **kw)

    @property
    def shape(self):
        return self.value.shape


def is_partitioned(x) -> bool:
    return isinstance(x, Partitioned)


def _is_leaf_default(x) -> bool:
    return not (isinstance(x, (dict, list, tuple)) or type(x) in _DATACLASSES) or x is None


def tree_map(fn: Callable, tree: Pytree, *rest: Pytree, is_leaf: Optional[Callable[[Any], bool]] = None) -> Pytree:
    """``jax.tree_util.tree_map`` for dict/list/tuple/registered dataclasses. ``None`` is a leaf-less node."""
    if is_leaf is not None and is_leaf(tree):
        return fn(tree, *rest)
    if tree is None:
        return None
    if isinstance(tree, dict):
        return type(tree)((k, tree_map(fn, v, *(r[k] for r in rest), is_leaf=is_leaf)) for k, v in tree.items())
    if isinstance(tree, tuple) and hasattr(tree, "_fields"):  # namedtuple
        return type(tree)(*(tree_map(fn, v, *(r[i] for r in rest), is_leaf=is_leaf) for i, v in enumerate(tree)))
    if isinstance(tree, (list, tuple)):
        return type(tree)(tree_map(fn, v, *(r[i] for r in rest), is_leaf=is_leaf) for i, v in enumerate(tree))
    if type(tree) in _DATACLASSES:
        data_f, _ = _DATACLASSES[type(tree)]
        upd = {f: tree_map(fn, getattr(tree, f), *(getattr(r, f) for r in rest), is_leaf=is_leaf) for f in data_f}
        return dataclasses.replace(tree, **upd)
    return fn(tree, *rest)


def tree_leaves(tree: Pytree, is_leaf: Optional[Callable[[Any], bool]] = None) -> list:
    out: list = []
    tree_map(lambda x: out.append(x), tree, is_leaf=is_leaf)
    return out


def tree_leaves_with_path(tree: Pytree, is_leaf=None, prefix: str = "") -> list[Tuple[str, Any]]:
    """Leaves with '/'-joined key paths (``input_dense/kernel``) — the checkpoint key format."""
    out: list[Tuple[str, Any]] = []

    def rec(t, path):
        if is_leaf is not None and is_leaf(t):
            out.append((path, t))
        elif t is None:
            return
        elif isinstance(t, dict):
            for k, v in t.items():
                rec(v, f"{path}/{k}" if path else str(k))
        elif isinstance(t, (list, tuple)):
            for i, v in enumerate(t):
                rec(v, f"{path}/{i}" if path else str(i))
        elif type(t) in _DATACLASSES:
            for f in _DATACLASSES[type(t)][0]:
                rec(getattr(t, f), f"{path}/{f}" if path else f)
        else:
            out.append((path, t))

    rec(tree, prefix)
    return out


def unbox(tree: Pytree) -> Pytree:
    """Strips ``Partitioned`` boxes, keeping values."""
    return tree_map(lambda x: x.value if is_partitioned(x) else x, tree, is_leaf=is_partitioned)


class PartitionSpec(tuple):
    """``P()`` = replicated, ``P("data")`` = dim 0 split over mesh axis "data" (jax.sharding.PartitionSpec)."""

    def __new__(cls, *names):
        return super().__new__(cls, names)

    def __reduce__(self):
        return (PartitionSpec, tuple(self))

    def __repr__(self):
        return "P(" + ", ".join(repr(n) for n in self) + ")"


P = PartitionSpec


def get_partition_spec(tree: Pytree) -> Pytree:
    """Spec pytree of a (boxed) state: ``Partitioned.names`` -> ``P(*names)``, everything else ``P()``.

    Reference: ``nn.get_partition_spec(state_shape)`` at param_sharding.py:266.
    """
    def spec(x):
        if is_partitioned(x):
            return P(*x.names)
        return P()
    return tree_map(spec, tree, is_leaf=is_partitioned)


def tree_zeros_like(tree: Pytree) -> Pytree:
    return tree_map(lambda x: torch.zeros_like(x) if isinstance(x, torch.Tensor) else type(x)(0), tree)


def tree_add(a: Pytree, b: Pytree) -> Pytree:
    return tree_map(lambda x, y: x + y, a, b)

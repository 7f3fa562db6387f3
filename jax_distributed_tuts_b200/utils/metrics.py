"""Console helpers (reference: util.py:12-14,170-185)."""
from __future__ import annotations

import textwrap

import torch

from ..tree import tree_leaves, unbox


def print_exception(e: BaseException) -> None:
    """Exception class name in red, message wrapped (util.py:12-14)."""
    name = f"\x1b[31m{type(e).__name__}\x1b[0m"
    print(textwrap.fill(f"{name}: {e}"))


def _host(x):
    return x.item() if isinstance(x, torch.Tensor) else x


def format_metrics(metrics, title: str = "") -> str:
    lines = [f"{k}: {_host(v[0]) / _host(v[1]):.6f}" for k, v in metrics.items()]
    if title:
        title = f" {title} "
        max_len = max(len(title), max(map(len, lines)))
        lines = [title.center(max_len, "=")] + lines
    return "\n".join(lines)


def print_metrics(metrics, title: str = "") -> None:
    """``name: sum/count`` with six decimals under a centred ``=`` title (util.py:170-181)."""
    print(format_metrics(metrics, title))


def get_num_params(state) -> int:
    """Total element count of ``state.params`` (util.py:184-185). For sharded states this counts
    the local slices, exactly as the reference would inside ``shard_map``. States owned by a fused
    engine are counted from the engine's layout (logical shapes, pads excluded)."""
    eng = getattr(state, "engine", None)
    if eng is not None:
        total = 0
        for sp in eng.layout.leaves.values():
            n = 1
            for s in sp.shape:
                n *= int(s)
            total += n
        return total
    total = 0
    for x in tree_leaves(unbox(state.params)):
        n = 1
        for s in x.shape:
            n *= int(s)
        total += n
    return total

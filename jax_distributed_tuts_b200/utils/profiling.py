"""Profiler scopes with the reference's region names (``jax.named_scope``: "sync_grads",
"sync_metrics", "shard_params", "gather_params", "sync_grad", "synch_metrics";
data_paral.py:210,220, param_sharding.py:58,146,350,355) emitted as NVTX ranges, plus CUDA-event
region timers (the reference has no timing code at all)."""
from __future__ import annotations

import contextlib
from collections import defaultdict
from typing import Dict, List

import torch

_TIMERS: Dict[str, List] = defaultdict(list)
_timing_enabled = False


def enable_region_timing(flag: bool = True) -> None:
    global _timing_enabled
    _timing_enabled = flag
    _TIMERS.clear()


def region_times_ms() -> Dict[str, float]:
    """Total device time per scope name since ``enable_region_timing`` (synchronises)."""
    out = {}
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    for name, pairs in _TIMERS.items():
        out[name] = sum(s.elapsed_time(e) for s, e in pairs)
    return out


class named_scope(contextlib.ContextDecorator):
    """Context manager *and* decorator, like ``jax.named_scope``."""

    def __init__(self, name: str):
        self.name = name
        self._pushed = False
        self._ev = None

    def __enter__(self):
        if torch.cuda.is_available() and torch.cuda.is_initialized():
            try:
                torch.cuda.nvtx.range_push(self.name)
                self._pushed = True
            except Exception:  # noqa: BLE001  (NVTX missing)
                self._pushed = False
            if _timing_enabled and not torch.cuda.is_current_stream_capturing():
                s = torch.cuda.Event(enable_timing=True)
                s.record()
                self._ev = s
        return self

    def __exit__(self, *exc):
        if self._ev is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            _TIMERS[self.name].append((self._ev, e))
            self._ev = None
        if self._pushed:
            torch.cuda.nvtx.range_pop()
            self._pushed = False
        return False

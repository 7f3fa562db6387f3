"""Learning-rate schedules in the shape optax gives them: a function ``count -> learning rate`` that can be
passed wherever a float learning rate is accepted by the generic-tier optimizers (``adamw``, ``sgd``).
The reference uses constant rates (data_paral.py:57-61, param_sharding.py:48); schedules are what a user of
``optax.adamw(learning_rate=schedule)`` would bring along. The fused engines bake a constant rate into their
captured step."""
from __future__ import annotations

import math
from typing import Callable

Schedule = Callable[[int], float]


def constant_schedule(value: float) -> Schedule:
    return lambda count: value


def linear_schedule(init_value: float, end_value: float, transition_steps: int, transition_begin: int = 0) -> Schedule:
    def fn(count: int) -> float:
        c = min(max(int(count) - transition_begin, 0), transition_steps)
        frac = c / transition_steps if transition_steps > 0 else 1.0
        return init_value + frac * (end_value - init_value)
    return fn


def cosine_decay_schedule(init_value: float, decay_steps: int, alpha: float = 0.0) -> Schedule:
    def fn(count: int) -> float:
        c = min(int(count), decay_steps)
        cos = 0.5 * (1.0 + math.cos(math.pi * c / max(decay_steps, 1)))
        return init_value * ((1.0 - alpha) * cos + alpha)
    return fn


def warmup_cosine_decay_schedule(init_value: float, peak_value: float, warmup_steps: int, decay_steps: int,
                                 end_value: float = 0.0) -> Schedule:
    """Linear warm-up from ``init_value`` to ``peak_value`` over ``warmup_steps``, then a cosine decay to
    ``end_value`` at ``decay_steps`` (total, warm-up included), as ``optax.warmup_cosine_decay_schedule``."""
    warm = linear_schedule(init_value, peak_value, warmup_steps)
    alpha = end_value / peak_value if peak_value else 0.0
    cos = cosine_decay_schedule(peak_value, max(decay_steps - warmup_steps, 1), alpha)
    return lambda count: warm(count) if int(count) < warmup_steps else cos(int(count) - warmup_steps)


def resolve(learning_rate, count) -> float:
    """The rate to use for the update that follows ``count`` completed updates."""
    return float(learning_rate(int(count))) if callable(learning_rate) else float(learning_rate)

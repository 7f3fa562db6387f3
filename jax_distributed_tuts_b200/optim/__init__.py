from .adamw import AdamW, EmptyState, ScaleByAdamState, adamw  # noqa: F401
from .sgd import SGD, TraceState, sgd  # noqa: F401
from .schedules import constant_schedule, cosine_decay_schedule, linear_schedule, warmup_cosine_decay_schedule  # noqa: F401

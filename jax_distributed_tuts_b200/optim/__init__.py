from .adamw import AdamW, EmptyState, ScaleByAdamState, adamw  # noqa: F401
from .sgd import SGD, TraceState, sgd  # noqa: F401

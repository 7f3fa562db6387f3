from .adamw import AdamW, EmptyState, ScaleByAdamState, adamw  # noqa: F401

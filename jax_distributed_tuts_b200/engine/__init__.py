from .base import FlatLayout, FusedEngine, SymmetricArena  # noqa: F401
from .gpt2 import GPT2Engine  # noqa: F401
from .mlp import MLPEngine  # noqa: F401

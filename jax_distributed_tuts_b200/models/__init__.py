from .mlp import Classifier, Dense, Module  # noqa: F401

"""B200-native multi-GPU training framework with the capabilities of jax-distributed-tuts."""
__version__ = "0.1.0"

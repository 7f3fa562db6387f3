"""Counter-based RNG keys (the analogue of ``jax.random.PRNGKey/split/fold_in``).

A :class:`Key` is a 64-bit state held in a one-element int64 tensor (``base``; on the GPU it
stays on the GPU so that CUDA graphs replay with fresh randomness) plus a host-side ``salt``
that accumulates ``split``/``fold_in`` constants without touching the device. The value a
kernel consumes is ``mix64(base ^ mix64(salt))``; the CUDA kernels and :func:`materialize`
compute exactly the same thing, and :func:`dropout_mask` reproduces the kernels' dropout mask
bit for bit in plain PyTorch (the CPU oracle for the fused GEMM epilogue).

Reference: ``jax.random.split`` / ``fold_in`` as used in data_paral.py:28-34,136,199 and
util.py:51. Threefry streams cannot be matched; the *structure* (per-step, per-minibatch,
per-device streams) is kept.
"""
from __future__ import annotations

from dataclasses import dataclass, replace
from typing import Tuple

import torch

MASK64 = (1 << 64) - 1
_GOLD = 0x9E3779B97F4A7C15


def mix64(z: int) -> int:
    """splitmix64 finaliser; must match ``b200::mix64`` in csrc/common.cuh."""
    z = (z + _GOLD) & MASK64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK64
    return z ^ (z >> 31)


def _to_signed(u: int) -> int:
    return u - (1 << 64) if u >= (1 << 63) else u


@dataclass(frozen=True)
class Key:
    base: torch.Tensor  # int64[1]
    salt: int = 0

    @property
    def device(self) -> torch.device:
        return self.base.device

    def to(self, device) -> "Key":
        return Key(self.base.to(device), self.salt)


def PRNGKey(seed: int, device: str | torch.device = "cpu") -> Key:
    return Key(torch.tensor([_to_signed(mix64(int(seed) & MASK64))], dtype=torch.int64, device=device), 0)


def fold_in(key: Key, data: int) -> Key:
    """Distinct stream per integer ``data`` (reference: data_paral.py:28-34)."""
    return replace(key, salt=mix64(key.salt ^ mix64((int(data) + 0x5BF03635) & MASK64)))


def split(key: Key, num: int = 2) -> Tuple[Key, ...]:
    return tuple(replace(key, salt=mix64(key.salt ^ mix64((i + (1 << 40)) & MASK64))) for i in range(num))


def salt_word(key: Key) -> int:
    """The host constant kernels fold into ``*base``: ``seed = mix64(base ^ salt_word)``."""
    return mix64(key.salt)


def materialize(key: Key) -> int:
    """Host value of the key's seed (synchronises if the base lives on the GPU)."""
    base = int(key.base.item()) & MASK64
    return mix64(base ^ salt_word(key))


def rebase(key: Key, inplace: bool = False) -> Key:
    """Folds the salt into the device state: ``Key(mix64(base ^ mix64(salt)), 0)``.

    One tiny kernel on the GPU; with ``inplace=True`` the same tensor is reused (what a CUDA
    graph replay needs: fixed addresses, evolving contents).
    """
    if key.base.is_cuda:
        from .ops import kernels as K
        out = key.base if inplace else torch.empty_like(key.base)
        if not inplace:
            out.copy_(key.base)
        K.rng_split_(out, None, salt_word(key), 0)
        return Key(out, 0)
    val = _to_signed(materialize(key))
    if inplace:
        key.base[0] = val
        return Key(key.base, 0)
    return Key(torch.tensor([val], dtype=torch.int64), 0)


def _hash32(x: torch.Tensor) -> torch.Tensor:
    m = 0xFFFFFFFF
    x = x & m
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & m
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & m
    x = x ^ (x >> 16)
    return x


def dropout_mask(seed: int, rows: int, cols: int, rate: float, device="cpu") -> torch.Tensor:
    """Boolean keep-mask identical to ``b200::dropout_keep4`` over a [rows, cols] matrix: three
    32-bit hashes of (seed, row, col // 4) give the four 16-bit uniforms of a 4-column group."""
    m32 = 0xFFFFFFFF
    thresh = int((1.0 - rate) * 65536.0 + 0.5)
    ngroups = (cols + 3) // 4
    r = torch.arange(rows, dtype=torch.int64, device=device)[:, None]
    g = torch.arange(ngroups, dtype=torch.int64, device=device)[None, :]
    s_lo, s_hi = seed & m32, (seed >> 32) & m32
    k0 = _hash32(s_lo ^ ((r * 0x9E3779B1) & m32) ^ ((g * 0x85EBCA77) & m32))
    k1 = _hash32(k0 ^ s_hi)
    k2 = _hash32((k1 + 0x6C8E9CF5) & m32)
    u = torch.stack([k1 & 0xFFFF, k1 >> 16, k2 & 0xFFFF, k2 >> 16], dim=-1)   # [rows, ngroups, 4]
    return (u < thresh).reshape(rows, ngroups * 4)[:, :cols]


def uniform_generator(key: Key) -> torch.Generator:
    """A torch CPU generator seeded from the key (parameter init, synthetic data)."""
    g = torch.Generator(device="cpu")
    g.manual_seed(materialize(key) & ((1 << 63) - 1))
    return g

"""Fused (sm_100a) backend of the named-axis collectives: psum / pmean / psum_scatter /
all_gather as ONE kernel launch each over NVSwitch, without NCCL.

Every mesh axis (or axis subset) gets a *symmetric workspace*: the same allocation on every rank
of the axis, mapped into every peer (``torch.distributed._symmetric_memory``), with an NVLS
multicast address when the fabric has one, plus a signal buffer for the epoch-flag protocol of
``csrc/comm.cuh``. A collective copies its operand into the workspace (a local device copy), runs
``b200_fused_collective`` (flag exchange -> ``multimem.ld_reduce`` of the rank's slice ->
``multimem.st`` / local store -> flag exchange) and hands back a tensor. Tensors larger than the
workspace go through it in chunks.

Replaces, for GPU runs of the generic tier, the NCCL thunks XLA would emit for
``jax.lax.psum/pmean`` (data_paral.py:212,223), ``psum_scatter`` (param_sharding.py:136) and
``all_gather`` (param_sharding.py:140). The fused *optimizer* paths (reduce + AdamW + publish in
one kernel) live in ``engine/``; this module is the same transport without the optimizer, for
code that calls collectives by axis name.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional, Tuple

import torch

from ..mesh import AxisNames, Mesh
from ..ops import _lib

_WORKSPACE_BYTES = int(os.environ.get("B200DIST_FUSED_WORKSPACE_MB", "64")) << 20
_GROUPS: Dict[Tuple[int, Tuple[str, ...]], "AxisWorkspace"] = {}


def available(mesh: Mesh) -> bool:
    """The fused backend needs GPUs (one per rank), the kernel library and an NCCL process group
    (symmetric memory rendezvous rides on it)."""
    import torch.distributed as dist
    return bool(mesh.device.type == "cuda" and _lib.available() and dist.is_available() and dist.is_initialized()
                and dist.get_backend() == "nccl")


class AxisWorkspace:
    """Symmetric buffer + signal pad shared by the ranks of one mesh axis."""

    def __init__(self, mesh: Mesh, axes: AxisNames, nbytes: int = _WORKSPACE_BYTES):
        from ..engine.base import SymmetricArena
        self.mesh, self.axes = mesh, axes
        self.world, self.rank = mesh.axis_size(axes), mesh.axis_index(axes)
        # floats per call: a multiple of 4 * world so every rank's slice is made of whole 16-byte vectors
        unit = 4 * self.world
        self.capacity = (nbytes // 4) // unit * unit
        self.arena = SymmetricArena({"buf": self.capacity * 4, "signal": _lib.SIGNAL_WORDS * 4}, mesh, mesh.device, axes=axes)
        self.buf = self.arena.tensor("buf", torch.float32, self.capacity)
        self._peer_buf = self.arena.peer_ptr_array("buf")
        self._peer_sig = self.arena.peer_ptr_array("signal")
        self.blocks = 64

    def _launch(self, mode: int, n: int, out: Optional[torch.Tensor], scale: float) -> None:
        a = _lib.CollectiveArgs()
        a.rank, a.world, a.mode, a.blocks = self.rank, self.world, mode, self.blocks
        a.peer_buf = C.cast(self._peer_buf, C.POINTER(C.c_void_p))
        a.peer_signal = C.cast(self._peer_sig, C.POINTER(C.c_void_p))
        a.mc_buf = self.arena.mc_ptr("buf") if mode != 2 else None
        a.out = out.data_ptr() if out is not None else None
        a.n, a.scale = n, scale
        _lib.check(_lib.lib().b200_fused_collective(C.byref(a), torch.cuda.current_stream().cuda_stream))

    # ---- collectives on flat fp32 data -----------------------------------------------------------
    def all_reduce_(self, flat: torch.Tensor, scale: float) -> None:
        """``flat`` (fp32, contiguous, any length) <- scale * sum over the axis, in place."""
        unit = 4 * self.world
        for lo in range(0, flat.numel(), self.capacity):
            hi = min(flat.numel(), lo + self.capacity)
            n = hi - lo
            npad = (n + unit - 1) // unit * unit
            self.buf[:n].copy_(flat[lo:hi])
            if npad > n:
                self.buf[n:npad].zero_()
            self._launch(0, npad, None, scale)
            flat[lo:hi].copy_(self.buf[:n])

    def reduce_scatter(self, chunks: torch.Tensor, scale: float) -> torch.Tensor:
        """``chunks`` [world, m] fp32 -> [m]: scale * sum over the axis of this rank's row."""
        W, m = chunks.shape
        assert W == self.world
        out = torch.empty(m, dtype=torch.float32, device=chunks.device)
        per_call = self.capacity // W // 4 * 4       # elements of one row per launch
        for lo in range(0, m, per_call):
            hi = min(m, lo + per_call)
            w = hi - lo
            wpad = (w + 3) // 4 * 4
            stage = self.buf[: W * wpad].view(W, wpad)
            stage[:, :w].copy_(chunks[:, lo:hi])
            if wpad > w:
                stage[:, w:].zero_()
            if wpad == w and hi == m and lo == 0:
                self._launch(1, W * wpad, out, scale)
            else:
                tmp = torch.empty(wpad, dtype=torch.float32, device=chunks.device)
                self._launch(1, W * wpad, tmp, scale)
                out[lo:hi].copy_(tmp[:w])
        return out

    def all_gather(self, x: torch.Tensor) -> torch.Tensor:
        """``x`` (contiguous, any dtype) -> [world, *x.shape]."""
        raw = x.reshape(-1).view(torch.uint8)
        nbytes = raw.numel()
        out = torch.empty((self.world, nbytes), dtype=torch.uint8, device=x.device)
        cap = self.capacity * 4 // 16 * 16
        stage_u8 = self.buf.view(torch.uint8)
        for lo in range(0, nbytes, cap):
            hi = min(nbytes, lo + cap)
            n = hi - lo
            npad = (n + 15) // 16 * 16
            stage_u8[:n].copy_(raw[lo:hi])
            if npad == n and lo == 0 and hi == nbytes:
                self._launch(2, npad // 16, out.view(-1), 1.0)
            else:
                tmp = torch.empty((self.world, npad), dtype=torch.uint8, device=x.device)
                self._launch(2, npad // 16, tmp.view(-1), 1.0)
                out[:, lo:hi].copy_(tmp[:, :n])
        return out.view(x.dtype).reshape((self.world,) + tuple(x.shape))


def reset() -> None:
    """Drops every axis workspace (their symmetric allocations are freed with them)."""
    _GROUPS.clear()


def workspace(mesh: Mesh, axes: AxisNames) -> AxisWorkspace:
    """The (lazily created) workspace of an axis. Collective on first use: every rank of the axis
    must make its first fused call on that axis at the same point of the program."""
    key = (id(mesh), mesh._norm(axes))
    ws = _GROUPS.get(key)
    if ws is None:
        ws = AxisWorkspace(mesh, axes)
        _GROUPS[key] = ws
    return ws


def _reducible(t: torch.Tensor) -> bool:
    return t.is_cuda and t.numel() > 0 and (t.dtype in (torch.float32, torch.bfloat16, torch.float16) or
                          (not t.dtype.is_floating_point and not t.dtype.is_complex and t.numel() <= 64))


def psum(t: torch.Tensor, axes: AxisNames, mesh: Mesh, mean: bool) -> Optional[torch.Tensor]:
    """Fused all-reduce of one tensor; ``None`` when the tensor is not something the kernel handles
    (the caller falls back to torch.distributed). Small integer tensors (metric counts) travel as
    fp32, which is exact below 2**24."""
    if not _reducible(t):
        return None
    ws = workspace(mesh, axes)
    if not t.dtype.is_floating_point and t.abs().max().item() * ws.world >= 2 ** 24:
        return None
    flat = t.detach().to(torch.float32).contiguous().reshape(-1).clone()
    ws.all_reduce_(flat, 1.0 / ws.world if mean else 1.0)
    out = flat.reshape(t.shape)
    if t.dtype.is_floating_point:
        return out.to(t.dtype)
    return out.round().to(t.dtype) if not mean else out


def psum_scatter(chunks: torch.Tensor, axes: AxisNames, mesh: Mesh) -> Optional[torch.Tensor]:
    """``chunks`` [world, ...] -> this rank's summed chunk."""
    if not (chunks.is_cuda and chunks.dtype in (torch.float32, torch.bfloat16, torch.float16)):
        return None
    ws = workspace(mesh, axes)
    W = chunks.shape[0]
    out = ws.reduce_scatter(chunks.to(torch.float32).reshape(W, -1), 1.0)
    return out.reshape(chunks.shape[1:]).to(chunks.dtype)


def all_gather(x: torch.Tensor, axes: AxisNames, mesh: Mesh) -> Optional[torch.Tensor]:
    if not x.is_cuda or x.numel() == 0:
        return None
    return workspace(mesh, axes).all_gather(x.contiguous())

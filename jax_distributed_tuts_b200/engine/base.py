"""Fused training engine: flat parameter/gradient buffers in symmetric memory, a hand-scheduled
forward/backward made of our sm_100a kernels, the fused collective+optimizer kernel, and CUDA
graph capture of the whole step.

This is the B200 counterpart of ``jax.jit(shard_map(train_step, ...), donate_argnames=...)``
(data_paral.py:241-251, param_sharding.py:370-379): the reference relies on a tracing compiler
to fuse and schedule; here the schedule is explicit, captured once with
``torch.cuda.CUDAGraph`` and replayed. State is updated in place (donation).
"""
from __future__ import annotations

import ctypes as C
import dataclasses
import os
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from .. import rng as _rng
from ..mesh import Mesh
from ..ops import _lib
from ..ops import kernels as K
from ..ops._lib import AdamWHParams, CommCtx, Seg
from ..state import Batch, Metrics, TrainState
from ..tree import Partitioned, Pytree


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclasses.dataclass
class LeafSpec:
    path: str                 # "input_dense/kernel"
    shape: Tuple[int, ...]    # logical shape
    rows: int
    cols: int
    ld: int                   # padded row pitch (multiple of 8)
    offset: int               # element offset in the flat buffers
    bucket: int


class FlatLayout:
    """Maps a parameter pytree onto one flat, padded buffer.

    Each leaf is stored row-major with its last dimension padded to a multiple of 8 elements
    (so bf16 rows are 16-byte multiples, which TMA requires) at an offset aligned to 64 elements.
    Leaves are grouped into buckets (gradient-ready order); every bucket is padded to a multiple
    of ``world * 64`` elements so a data-parallel rank's 1/world slice is vector-aligned.
    The same layout is used for fp32 master params, fp32 grads, bf16 working copy, m and v, so
    the optimizer and the collectives are pure flat-buffer kernels; pad elements stay zero.
    """

    def __init__(self, leaves: Sequence[Tuple[str, Tuple[int, ...], int]], world: int):
        self.world = world
        self.leaves: Dict[str, LeafSpec] = {}
        self.bucket_ranges: List[Tuple[int, int]] = []
        off = 0
        cur_bucket, bucket_start = None, 0
        for path, shape, bucket in leaves:
            if cur_bucket is None:
                cur_bucket = bucket
            if bucket != cur_bucket:
                off = _round_up(off, world * 64)
                self.bucket_ranges.append((bucket_start, off))
                bucket_start, cur_bucket = off, bucket
            cols = shape[-1]
            rows = 1
            for s in shape[:-1]:
                rows *= s
            ld = _round_up(cols, 8)
            self.leaves[path] = LeafSpec(path, tuple(shape), rows, cols, ld, off, bucket)
            off = _round_up(off + rows * ld, 64)
        off = _round_up(off, world * 64)
        self.bucket_ranges.append((bucket_start, off))
        self.total = off

    def view(self, buf: torch.Tensor, path: str) -> torch.Tensor:
        """Logical-shape view of one leaf inside a flat buffer."""
        sp = self.leaves[path]
        v = buf[sp.offset: sp.offset + sp.rows * sp.ld].view(sp.rows, sp.ld)[:, : sp.cols]
        if len(sp.shape) == 1:
            return v[0]
        return v.view(sp.shape) if sp.ld == sp.cols else v

    def view2d(self, buf: torch.Tensor, path: str) -> torch.Tensor:
        sp = self.leaves[path]
        return buf[sp.offset: sp.offset + sp.rows * sp.ld].view(sp.rows, sp.ld)[:, : sp.cols]

    def tree(self, buf: torch.Tensor) -> Dict:
        out: Dict = {}
        for path in self.leaves:
            node = out
            *parents, last = path.split("/")
            for p in parents:
                node = node.setdefault(p, {})
            node[last] = self.view(buf, path)
        return out


class SymmetricArena:
    """One symmetric-memory allocation carved into named sub-buffers.

    With world > 1 the arena comes from ``torch.distributed._symmetric_memory`` (VMM allocation
    mapped into every rank, NVLS multicast address when the fabric supports it); with world == 1
    it is an ordinary device allocation. Sub-buffers are 1 KiB aligned."""

    def __init__(self, sizes_bytes: Dict[str, int], mesh: Mesh, device: torch.device):
        self.world, self.rank = mesh.size, mesh.rank
        self.offsets: Dict[str, int] = {}
        off = 0
        for name, nbytes in sizes_bytes.items():
            self.offsets[name] = off
            off = _round_up(off + nbytes, 1024)
        self.nbytes = off
        self.multicast_ptr = 0
        self.backend = "local"
        if self.world > 1:
            import torch.distributed._symmetric_memory as symm_mem
            gname = dist.group.WORLD.group_name
            self.buf = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=device)
            self.buf.zero_()
            hdl = symm_mem.rendezvous(self.buf, gname)
            self.handle = hdl
            self.peer_ptrs = [int(p) for p in hdl.buffer_ptrs]
            try:
                self.multicast_ptr = int(hdl.multicast_ptr) if os.environ.get("B200DIST_NO_MULTICAST", "0") != "1" else 0
            except Exception:  # noqa: BLE001
                self.multicast_ptr = 0
            self.backend = "symm_mem+nvls" if self.multicast_ptr else "symm_mem+p2p"
            torch.cuda.synchronize()
            dist.barrier()
        else:
            self.buf = torch.zeros(self.nbytes, dtype=torch.uint8, device=device)
            self.peer_ptrs = [self.buf.data_ptr()]

    def tensor(self, name: str, dtype: torch.dtype, numel: int) -> torch.Tensor:
        o = self.offsets[name]
        return self.buf[o: o + numel * torch.empty(0, dtype=dtype).element_size()].view(dtype)

    def peer_ptr_array(self, name: str):
        arr = (C.c_void_p * self.world)()
        for i, p in enumerate(self.peer_ptrs):
            arr[i] = p + self.offsets[name]
        return arr

    def mc_ptr(self, name: str) -> Optional[int]:
        return self.multicast_ptr + self.offsets[name] if self.multicast_ptr else None


class FusedEngine:
    """Common machinery; subclasses provide the layout and the per-minibatch forward/backward."""

    metric_names = ("loss", "accuracy")

    def __init__(self, mesh: Mesh, *, lr: float, b1: float = 0.9, b2: float = 0.999, eps: float = 1e-8,
                 weight_decay: float = 1e-4, num_minibatches: int = 4, data_axis_name: str = "data",
                 use_graph: bool = True, comm_blocks: int = 16):
        if not torch.cuda.is_available():
            raise RuntimeError("the fused engine needs a CUDA device; use the generic tier on CPU")
        _lib.lib()  # fail loudly if the kernel library is missing
        self.mesh = mesh
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.world, self.rank = mesh.size, mesh.rank
        self.hp = dict(lr=lr, beta1=b1, beta2=b2, eps=eps, weight_decay=weight_decay)
        self.num_minibatches = num_minibatches
        self.data_axis_name = data_axis_name
        self.use_graph = use_graph
        self.comm_blocks = comm_blocks
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self._last_metrics = None
        self._counts = [0, 0]
        self.side_stream = torch.cuda.Stream()
        self.steps_done = 0

    # -- buffers shared by all engines -------------------------------------------------------
    def _alloc_common(self) -> None:
        d = self.device
        self.step_count = torch.zeros(1, dtype=torch.int32, device=d)
        self.step_metrics = torch.zeros(8, dtype=torch.float32, device=d)
        self.metrics_acc = torch.zeros(8, dtype=torch.float32, device=d)

    def _make_ctx(self, arena: SymmetricArena, grad: str, param: Optional[str], bf16: Optional[str],
                  signal: str) -> CommCtx:
        ctx = CommCtx()
        ctx.rank, ctx.world = self.rank, self.world
        self._keep = getattr(self, "_keep", [])
        def arr(name):
            a = arena.peer_ptr_array(name)
            self._keep.append(a)
            return C.cast(a, C.POINTER(C.c_void_p))
        ctx.peer_grad = arr(grad)
        if param:
            ctx.peer_param = arr(param)
            ctx.mc_param = arena.mc_ptr(param)
        if bf16:
            ctx.peer_bf16 = arr(bf16)
            ctx.mc_bf16 = arena.mc_ptr(bf16)
        ctx.peer_signal = arr(signal)
        ctx.mc_grad = arena.mc_ptr(grad)
        return ctx

    def _segs_to_device(self, segs: List[Seg]) -> torch.Tensor:
        raw = b"".join(bytes(s) for s in segs)
        t = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(self.device)
        return t

    def _fused_update(self, ctx: CommCtx, segs_dev: torch.Tensor, nseg: int, *, broadcast: bool,
                      p_local, bf16_local, m, v, grad_scale: float, bump: bool, with_metrics: bool) -> None:
        hp = AdamWHParams(self.hp["lr"], self.hp["beta1"], self.hp["beta2"], self.hp["eps"],
                          self.hp["weight_decay"], grad_scale)
        nmet = len(self.metric_names)
        _lib.check(_lib.lib().b200_fused_reduce_adamw(
            C.byref(ctx), segs_dev.data_ptr(), nseg, int(broadcast),
            None if p_local is None else p_local.data_ptr(),
            None if bf16_local is None else bf16_local.data_ptr(),
            m.data_ptr(), v.data_ptr(), C.byref(hp), self.step_count.data_ptr(), int(bump),
            self.step_metrics.data_ptr() if with_metrics else None,
            self.metrics_acc.data_ptr() if with_metrics else None, nmet if with_metrics else 0,
            self.comm_blocks, torch.cuda.current_stream().cuda_stream))

    # -- step driver --------------------------------------------------------------------------
    def _load_batch(self, batch: Batch) -> None:  # pragma: no cover - interface
        raise NotImplementedError

    def _step_body(self) -> None:  # pragma: no cover - interface
        raise NotImplementedError

    def _run_step(self) -> None:
        """Call 1 runs the step eagerly (lazy initialisation inside CUDA/NCCL happens here), call 2
        captures it into a CUDA graph and replays it, later calls only replay. Every call performs
        exactly one real training step."""
        if not self.use_graph:
            before = _lib.launch_count()
            self._step_body()
            self.launches_per_step = _lib.launch_count() - before
            return
        if self.graph is None:
            if not getattr(self, "_warmed", False):
                s = torch.cuda.Stream()
                s.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(s):
                    before = _lib.launch_count()
                    self._step_body()
                    self.launches_per_step = _lib.launch_count() - before
                torch.cuda.current_stream().wait_stream(s)
                self._warmed = True
                return
            torch.cuda.synchronize()
            if self.world > 1:
                dist.barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=torch.cuda.Stream()):
                self._step_body()
            self.graph = g
        self.graph.replay()

    def train_step(self, state: TrainState, metrics: Optional[Metrics], batch: Batch) -> Tuple[TrainState, Metrics]:
        """``(state, metrics, batch) -> (state, metrics)`` like the reference's jitted steps. The
        state's buffers are updated in place; the returned state replaces the argument."""
        # running metrics: continue in place if the caller passes back what we returned last
        if metrics is None or metrics is not self._last_metrics:
            self.metrics_acc.zero_()
            self._counts = [0, 0]
            if metrics is not None:
                for i, name in enumerate(self.metric_names):
                    s, c = metrics[name]
                    self.metrics_acc[i] = float(s)
                    self._counts[i] = int(c)
        self._load_batch(batch)
        self._run_step()
        self.steps_done += 1
        per_step = self._rows_per_step()
        out = {}
        for i, name in enumerate(self.metric_names):
            self._counts[i] += per_step
            out[name] = (self.metrics_acc[i], self._counts[i])
        self._last_metrics = out
        new_state = dataclasses.replace(state, step=state.step + 1)
        return new_state, out

    def _rows_per_step(self) -> int:  # pragma: no cover - interface
        raise NotImplementedError

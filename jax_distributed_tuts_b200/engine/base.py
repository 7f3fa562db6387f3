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
from ..parallel.fsdp import choose_shard_dim
from ..tree import Partitioned


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

    # Owner-update data parallelism keeps per-rank state (Adam moments) only for the rank's 1/world
    # slice of every bucket; a rank's local buffer is the concatenation of its slices, bucket by bucket.
    def owner_slices_to_flat(self, slices: torch.Tensor) -> torch.Tensor:
        """``slices`` [world, total/world] (row r = rank r's local buffer) -> the flat [total] buffer."""
        W = self.world
        assert slices.shape == (W, self.total // W), (tuple(slices.shape), W, self.total)
        full = torch.empty(self.total, dtype=slices.dtype, device=slices.device)
        for (b0, b1) in self.bucket_ranges:
            sl = (b1 - b0) // W
            full[b0:b1].view(W, sl).copy_(slices[:, b0 // W: b0 // W + sl])
        return full

    def flat_to_owner_slice(self, full: torch.Tensor, rank: int) -> torch.Tensor:
        """Inverse of ``owner_slices_to_flat`` for one rank: flat [total] -> local [total/world]."""
        W = self.world
        out = torch.empty(self.total // W, dtype=full.dtype, device=full.device)
        for (b0, b1) in self.bucket_ranges:
            sl = (b1 - b0) // W
            out[b0 // W: b0 // W + sl].copy_(full[b0 + rank * sl: b0 + (rank + 1) * sl])
        return out

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

    With more than one rank the arena comes from ``torch.distributed._symmetric_memory`` (VMM
    allocation mapped into every rank of the group, NVLS multicast address when the fabric supports
    it); with one rank it is an ordinary device allocation. Sub-buffers are 1 KiB aligned.

    ``axes`` restricts the arena to the ranks of one mesh axis (or axis subset): the rendezvous runs
    on that axis' communicator, ``rank`` / ``world`` are the coordinates within it and the multicast
    address covers exactly those GPUs (a per-axis symmetric group for N-D meshes)."""

    def __init__(self, sizes_bytes: Dict[str, int], mesh: Mesh, device: torch.device, axes=None):
        if axes is None:
            self.world, self.rank = mesh.size, mesh.rank
            group = dist.group.WORLD if mesh.size > 1 else None
        else:
            self.world, self.rank = mesh.axis_size(axes), mesh.axis_index(axes)
            group = mesh.group(axes) if self.world > 1 else None
            if self.world > 1:
                ranks = mesh.group_ranks(axes)
                assert ranks.index(mesh.rank) == self.rank, (ranks, mesh.rank, self.rank)
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
            gname = group.group_name
            if hasattr(symm_mem, "enable_symm_mem_for_group"):
                import warnings
                with warnings.catch_warnings():      # (deprecated no-op in newer torch, required by older ones)
                    warnings.simplefilter("ignore")
                    try:
                        symm_mem.enable_symm_mem_for_group(gname)
                    except Exception:  # noqa: BLE001
                        pass
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
            dist.barrier(group=group)
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
    """Common machinery: flat symmetric buffers, shard planning (the reference's shard rule),
    weight access for the GEMMs (in place from peer shards in FSDP mode), the fused
    collective+optimizer launches, CUDA-graph capture, metrics and RNG bookkeeping.
    Subclasses provide the leaf list and the forward/backward schedule.

    Modes
      dp    parameters replicated; every rank reduces + updates 1/world of each bucket and
            multicasts the result (data_paral.py:210-217).
      fsdp  leaves sharded over ``shard_axis`` by the reference's rule (param_sharding.py:89-113);
            gradients are reduced over *all* ranks of the mesh (so on a 2-D ``(data, fsdp)`` mesh
            the data-axis mean of ``sync_gradients``, param_sharding.py:293-322, is the same
            launch), weights are read from the peers of the rank's shard group.
    """

    metric_names = ("loss", "accuracy")

    def __init__(self, mesh: Mesh, *, mode: str = "dp", lr: float, b1: float = 0.9, b2: float = 0.999,
                 eps: float = 1e-8, weight_decay: float = 1e-4, num_minibatches: int = 4,
                 data_axis_name: str = "data", shard_axis: Optional[str] = None, min_weight_size: int = 2 ** 18,
                 use_graph: bool = True, comm_blocks: int = 16, overlap: bool = True, fsdp_weights: str = "fused",
                 publish_fp32: bool = False, stub_collectives: bool = False, deterministic: bool = False,
                 comm_axes=None):
        if not torch.cuda.is_available():
            raise RuntimeError("the fused engine needs a CUDA device; use the generic tier on CPU")
        _lib.lib()  # fail loudly if the kernel library is missing
        assert mode in ("dp", "fsdp")
        self.mesh, self.mode = mesh, mode
        self.device = torch.device("cuda", torch.cuda.current_device())
        # comm_axes: the mesh axes the engine's gradients are reduced over (default: the whole mesh). A
        # pipeline stage passes ("data",): its replicas along the data axis form the symmetric group,
        # the pipe axis is somebody else's business.
        self.comm_axes = comm_axes
        if comm_axes is None:
            self.world, self.rank = mesh.size, mesh.rank
            self.group = dist.group.WORLD if mesh.size > 1 else None
        else:
            self.world, self.rank = mesh.axis_size(comm_axes), mesh.axis_index(comm_axes)
            self.group = mesh.group(comm_axes) if self.world > 1 else None
        self.hp = dict(lr=lr, beta1=b1, beta2=b2, eps=eps, weight_decay=weight_decay)
        self.num_minibatches = num_minibatches
        self.data_axis_name = data_axis_name
        self.shard_axis = shard_axis or data_axis_name
        self.S = mesh.axis_size(self.shard_axis) if mode == "fsdp" else 1
        self.sidx = mesh.axis_index(self.shard_axis) if mode == "fsdp" else 0
        self.shard_peer_ranks = list(mesh.group_ranks(self.shard_axis)) if (mode == "fsdp" and self.S > 1) else [self.rank]
        self.min_weight_size = min_weight_size
        self.use_graph = use_graph
        self.comm_blocks = int(os.environ.get("B200DIST_COMM_BLOCKS", comm_blocks))     # (environment: tuning sweeps)
        self.overlap = overlap and self.world > 1
        # FSDP weight access: "fused" = the GEMM's TMA producer reads the peer shards in place (best
        # when the GEMM has few row tiles: every tile of the weight crosses NVLink once per row tile);
        # "gather" = a layer's shards are assembled into a local buffer by peer loads once per use,
        # prefetched on the side stream, and the GEMMs run on local (L2-cached) memory.
        assert fsdp_weights in ("fused", "gather"), "engines resolve 'auto' before calling the base class"
        self.fsdp_weights = fsdp_weights if (mode == "fsdp" and self.S > 1) else "fused"
        # DP publishes the bf16 working copy to every rank; the fp32 master of a slice stays with its
        # owner (like the Adam moments) unless publish_fp32 asks for the reference's fully replicated
        # fp32 state (data_paral.py:214-217): one third of the publish traffic. Readers of the fp32
        # tree (flat_params, checkpoints) gather the owners' slices first.
        self.publish_fp32 = publish_fp32
        self._master_stale = False
        # Run-to-run bit reproducibility. The default schedules sum some things in arrival order:
        # split-K partial tiles (red.global.add in the wgrad GEMMs), bias-gradient column sums added
        # atomically from GEMM / cross-entropy / LayerNorm epilogues, the embedding gradient scatter,
        # attention's dQ tile reduce-adds, and the loss / accuracy metric sums. With deterministic=True an
        # engine that supports it switches every gradient-carrying one of those to a fixed order (metric
        # sums stay atomics: they do not feed the parameters).
        self.deterministic = deterministic
        # Measurement aid (bench.py "exposed comm"): the same step with every cross-GPU access removed.
        # Each rank updates its own part from its local gradients only; results are NOT a training run.
        self.stub_collectives = stub_collectives and self.world > 1
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self._last_metrics = None
        self._counts = [0, 0]
        self.side_stream = torch.cuda.Stream()
        self.steps_done = 0
        self._keep: list = []

    # ---------------------------------------------------------------- buffers -----------------
    def _setup_buffers(self, layout: FlatLayout) -> None:
        self.layout = layout
        d, n, W = self.device, layout.total, self.world
        self.step_count = torch.zeros(1, dtype=torch.int32, device=d)
        self.step_metrics = torch.zeros(8, dtype=torch.float32, device=d)
        self.metrics_acc = torch.zeros(8, dtype=torch.float32, device=d)
        sizes = {"grad": n * 4, "signal": _lib.SIGNAL_WORDS * 4}
        if self.mode == "dp":
            sizes.update({"param": n * 4, "bf16": n * 2})
        else:
            self._plan_shards()
            sizes.update({"bf16": self.local_total * 2})   # every rank's bf16 slices at the same local offsets
        self.arena = SymmetricArena(sizes, self.mesh, d, axes=self.comm_axes)
        self.g32 = self.arena.tensor("grad", torch.float32, n)
        if self.mode == "dp":
            self.p32 = self.arena.tensor("param", torch.float32, n)
            self.pb = self.arena.tensor("bf16", torch.bfloat16, n)
            self.m = torch.zeros(n // W, dtype=torch.float32, device=d)
            self.v = torch.zeros(n // W, dtype=torch.float32, device=d)
        else:
            self.p_loc = torch.zeros(self.local_total, dtype=torch.float32, device=d)
            self.m = torch.zeros(self.local_total, dtype=torch.float32, device=d)
            self.v = torch.zeros(self.local_total, dtype=torch.float32, device=d)
            self.pb = self.arena.tensor("bf16", torch.bfloat16, self.local_total)
            self._setup_bias_gather()
            if self.fsdp_weights == "gather":
                self.gather_stream = torch.cuda.Stream()
                self._lru = 0
                self._setup_bucket_gather()
        self.ctx = self._make_ctx(self.arena, "grad", "param" if self.mode == "dp" else None, "bf16", "signal")
        if self.stub_collectives:
            self.ctx_stub = self._make_ctx(self.arena, "grad", "param" if self.mode == "dp" else None, "bf16", "signal",
                                           local_only=True)
        self._build_segments()

    def _plan_shards(self) -> None:
        self.shard_dim: Dict[str, Optional[int]] = {}
        self.loc_off: Dict[str, int] = {}
        loc = 0
        for path, sp in self.layout.leaves.items():
            numel = 1
            for s_ in sp.shape:
                numel *= s_
            dim = None
            if numel > self.min_weight_size:
                dim = choose_shard_dim(sp.shape, (None,) * len(sp.shape), self.S)
                dim = self._override_shard_dim(path, sp.shape, dim)
            self.shard_dim[path] = dim
            self.loc_off[path] = loc
            loc = _round_up(loc + self._local_numel(path), 64)
        self.local_total = _round_up(loc, 64)

    def _override_shard_dim(self, path: str, shape, dim):
        """Hook for leaves with access patterns the generic rule does not know about."""
        return dim

    def _local_shape2d(self, path: str) -> Tuple[int, int, int]:
        """(rows, cols, ld) of this rank's slice of a leaf in padded 2-D form."""
        sp = self.layout.leaves[path]
        dim = self.shard_dim[path]
        if dim is None:
            return sp.rows, sp.cols, sp.ld
        if len(sp.shape) == 1 or dim == len(sp.shape) - 1:
            cols = sp.cols // self.S
            return sp.rows, cols, _round_up(cols, 8)
        assert dim == 0, "leaves are sharded on their first or last dimension"
        return sp.rows // self.S, sp.cols, sp.ld

    def _local_numel(self, path: str) -> int:
        r, _, ld = self._local_shape2d(path)
        return r * ld

    def _local_view(self, buf: torch.Tensor, path: str) -> torch.Tensor:
        r, c, ld = self._local_shape2d(path)
        o = self.loc_off[path]
        return buf[o: o + r * ld].view(r, ld)[:, :c]

    def _setup_bias_gather(self) -> None:
        """Sharded 1-D leaves (biases, norm scales) are re-assembled once per step by peer loads."""
        from ..ops._lib import GatherSeg
        self.bias_full: Dict[str, torch.Tensor] = {}
        segs, off, spans = [], 0, {}
        for path, sp in self.layout.leaves.items():
            if len(sp.shape) == 1 and self.shard_dim[path] is not None:
                per = sp.cols // self.S
                segs.append(GatherSeg(self.loc_off[path] * 2, off, per * 2, 1, per * 2, per * 2, per * 2))
                spans[path] = (off // 2, sp.cols)
                off = _round_up(off + sp.cols * 2, 64)
        self._bias_nseg = len(segs)
        if segs:
            self._bias_buf = torch.zeros(off // 2, dtype=torch.bfloat16, device=self.device)
            self._bias_segs = self._segs_to_device(segs)
            arr = (C.c_void_p * self.S)()
            for i, r in enumerate(self.shard_peer_ranks):
                arr[i] = self.arena.peer_ptrs[r] + self.arena.offsets["bf16"]
            self._bias_peers = arr
            self._bias_peers_stub = (C.c_void_p * self.S)(*[self.arena.peer_ptrs[self.rank] + self.arena.offsets["bf16"]] * self.S)
            self.bias_full = {k: self._bias_buf[o: o + n_] for k, (o, n_) in spans.items()}

    def _setup_bucket_gather(self) -> None:
        """"gather" mode: per bucket, segment tables that assemble every sharded 2-D leaf of the bucket
        into one of two rotating local buffers laid out like the DP working copy."""
        from ..ops._lib import GatherSeg
        self._gbuf_segs: Dict[int, Tuple[torch.Tensor, int]] = {}
        per_bucket: Dict[int, list] = {}
        size = 0
        for (b0, b1) in self.layout.bucket_ranges:
            size = max(size, b1 - b0)
        for path, sp in self.layout.leaves.items():
            dim = self.shard_dim[path]
            if dim is None or len(sp.shape) == 1:
                continue
            r, c, ld = self._local_shape2d(path)
            base = self.layout.bucket_ranges[sp.bucket][0]
            dst = (sp.offset - base) * 2
            if c != sp.cols:     # column shards: peers' column blocks interleave into the full rows
                seg = GatherSeg(self.loc_off[path] * 2, dst, c * 2, sp.rows, ld * 2, sp.ld * 2, c * 2)
            else:                # row shards: contiguous blocks
                seg = GatherSeg(self.loc_off[path] * 2, dst, r * sp.ld * 2, 1, r * sp.ld * 2, r * sp.ld * 2, r * sp.ld * 2)
            per_bucket.setdefault(sp.bucket, []).append(seg)
        for b, segs in per_bucket.items():
            self._gbuf_segs[b] = (self._segs_to_device(segs), len(segs))
        self._gbufs = [torch.zeros(size, dtype=torch.bfloat16, device=self.device) for _ in range(2)]
        self._gbuf_bucket = [None, None]
        self._gbuf_events = [None, None]
        arr = (C.c_void_p * self.S)()
        for i, r_ in enumerate(self.shard_peer_ranks):
            arr[i] = self.arena.peer_ptrs[r_] + self.arena.offsets["bf16"]
        self._gather_peers = arr
        self._gather_peers_stub = (C.c_void_p * self.S)(*[self.arena.peer_ptrs[self.rank] + self.arena.offsets["bf16"]] * self.S)

    def prefetch_bucket(self, b: int, keep: Optional[int] = None) -> None:
        """Starts assembling bucket ``b``'s weights on the gather stream (no-op unless in gather
        mode). ``keep`` names the bucket the caller is about to use: its buffer is not recycled."""
        if self.fsdp_weights != "gather" or b not in self._gbuf_segs or b in self._gbuf_bucket:
            return
        if None in self._gbuf_bucket:
            slot = self._gbuf_bucket.index(None)
        else:
            cands = [i for i in (0, 1) if self._gbuf_bucket[i] != keep]
            slot = self._lru if self._lru in cands else cands[0]
        main = torch.cuda.current_stream()
        self.gather_stream.wait_stream(main)      # every reader of the recycled buffer is already queued on main
        segs, nseg = self._gbuf_segs[b]
        with torch.cuda.stream(self.gather_stream):
            _lib.check(_lib.lib().b200_peer_gather(C.cast(self._gather_peers_stub if self.stub_collectives else self._gather_peers,
                                                           C.POINTER(C.c_void_p)), self.S,
                                                    segs.data_ptr(), nseg, self._gbufs[slot].data_ptr(), 32,
                                                    self.gather_stream.cuda_stream))
            ev = torch.cuda.Event()
            ev.record(self.gather_stream)
        self._gbuf_bucket[slot] = b
        self._gbuf_events[slot] = ev
        self._lru = 1 - slot

    def _gathered(self, path: str) -> torch.Tensor:
        sp = self.layout.leaves[path]
        b = sp.bucket
        if b not in self._gbuf_bucket:
            self.prefetch_bucket(b, keep=b)
        slot = self._gbuf_bucket.index(b)
        ev = self._gbuf_events[slot]
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)
            self._gbuf_events[slot] = None
        base = self.layout.bucket_ranges[b][0]
        o = sp.offset - base
        return self._gbufs[slot][o: o + sp.rows * sp.ld].view(sp.rows, sp.ld)[:, : sp.cols]

    def release_gathered(self) -> None:
        """Forget which buckets the rotating buffers hold (call at step boundaries: the shards change)."""
        if self.fsdp_weights == "gather":
            self._gbuf_bucket = [None, None]
            self._gbuf_events = [None, None]
            self._lru = 0

    def _gather_small(self) -> None:
        if self.mode != "fsdp" or not self._bias_nseg:
            return
        _lib.check(_lib.lib().b200_peer_gather(C.cast(self._bias_peers_stub if self.stub_collectives else self._bias_peers,
                                                       C.POINTER(C.c_void_p)), self.S,
                                                self._bias_segs.data_ptr(), self._bias_nseg,
                                                self._bias_buf.data_ptr(), 2,
                                                torch.cuda.current_stream().cuda_stream))

    # ---------------------------------------------------------------- parameters --------------
    def load_flat_params(self, flat: Dict[str, torch.Tensor]) -> None:
        """Loads *global* (unsharded) parameter values, path -> tensor; every rank passes the same
        values (the reference initialises the full parameter on every device from one rng and keeps
        its slice, param_sharding.py:227-241)."""
        for path, sp in self.layout.leaves.items():
            v = flat[path]
            v = torch.as_tensor(v.value if isinstance(v, Partitioned) else v).to(self.device, dtype=torch.float32)   # tensors or numpy (checkpoints)
            v2 = v.reshape(sp.rows, sp.cols)
            if self.mode == "dp":
                self.layout.view2d(self.p32, path).copy_(v2)
                continue
            dim = self.shard_dim[path]
            r, c, ld = self._local_shape2d(path)
            if dim is not None:
                v2 = v2[:, self.sidx * c:(self.sidx + 1) * c] if c != sp.cols else v2[self.sidx * r:(self.sidx + 1) * r]
            self._local_view(self.p_loc, path).copy_(v2)
        if self.mode == "dp":
            K.cast_f32_bf16(self.p32, self.pb)
        else:
            K.cast_f32_bf16(self.p_loc, self.pb)
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier(group=self.group)

    def flat_params(self, gather: bool = True) -> Dict[str, object]:
        """path -> fp32 tensor. DP: views of the replicated master. FSDP: full arrays gathered over
        the shard axis (``gather=True``) or ``Partitioned`` boxes of the local slices."""
        out: Dict[str, object] = {}
        if self.mode == "dp":
            self._refresh_master()
        for path, sp in self.layout.leaves.items():
            if self.mode == "dp":
                out[path] = self.layout.view(self.p32, path)
                continue
            loc = self._local_view(self.p_loc, path)          # [rows, cols] view of the local slice
            dim = self.shard_dim[path]
            if len(sp.shape) == 1:
                loc = loc.reshape(-1)
            if dim is None:
                out[path] = loc.reshape(sp.shape)
            elif gather:
                from ..comm import collectives as Cc
                full = Cc.all_gather(loc.contiguous(), self.shard_axis, axis=dim, mesh=self.mesh) if self.S > 1 else loc
                out[path] = full.reshape(sp.shape)
            else:
                names = tuple(self.shard_axis if i == dim else None for i in range(len(sp.shape)))
                out[path] = Partitioned(loc, names)
        return out

    def _refresh_master(self) -> None:
        """DP without ``publish_fp32``: a rank's fp32 master is current only on the slices it owns.
        All-gathers the owners' slices bucket by bucket so ``p32`` is the full replicated master again.
        Collective (every rank calls ``flat_params`` / ``params_tree`` / checkpoint save together)."""
        if self.mode != "dp" or self.world == 1 or self.publish_fp32 or not self._master_stale:
            return
        torch.cuda.synchronize()
        W, r = self.world, self.rank
        for (b0, b1) in self.layout.bucket_ranges:
            sl = (b1 - b0) // W
            dist.all_gather_into_tensor(self.p32[b0:b1], self.p32[b0 + r * sl: b0 + (r + 1) * sl].clone(), group=self.group)
        torch.cuda.synchronize()
        self._master_stale = False

    # ---- optimizer state in the reference's (global, per-leaf) layout: checkpoints, inspection ----
    def _moment_flat(self, buf: torch.Tensor) -> Dict[str, torch.Tensor]:
        """path -> full fp32 array of an Adam moment (``buf`` is ``self.m`` or ``self.v``). DP keeps a
        moment only for the rank's 1/world slice of every bucket (owner-update): the slices are
        all-gathered and put back at their flat offsets. FSDP keeps it per local shard: sharded
        leaves are all-gathered over the shard axis like the parameters. Collective: every rank calls."""
        out: Dict[str, torch.Tensor] = {}
        if self.mode == "dp":
            W = self.world
            if W == 1:
                full = buf
            else:
                g = torch.empty(W, buf.numel(), dtype=buf.dtype, device=buf.device)
                dist.all_gather_into_tensor(g.view(-1), buf.contiguous().view(-1), group=self.group)
                full = self.layout.owner_slices_to_flat(g)
            for path, sp in self.layout.leaves.items():
                out[path] = self.layout.view(full, path).clone().reshape(sp.shape)
            return out
        from ..comm import collectives as Cc
        for path, sp in self.layout.leaves.items():
            loc = self._local_view(buf, path)
            dim = self.shard_dim[path]
            if len(sp.shape) == 1:
                loc = loc.reshape(-1)
            if dim is not None and self.S > 1:
                loc = Cc.all_gather(loc.contiguous(), self.shard_axis, axis=dim, mesh=self.mesh)
            out[path] = loc.reshape(sp.shape).clone()
        return out

    def opt_state_global(self) -> Dict[str, object]:
        """``{"count", "mu", "nu"}`` with ``mu`` / ``nu`` shaped like the parameter tree (what
        ``optax.adamw``'s ScaleByAdamState holds in the reference, data_paral.py:106-108), whatever
        the engine's mode and world size. Collective."""
        return {"count": self.step_count.clone(), "mu": self._nest(self._moment_flat(self.m)),
                "nu": self._nest(self._moment_flat(self.v))}

    def _load_moment(self, buf: torch.Tensor, flat: Dict[str, object]) -> None:
        if self.mode == "dp":
            full = torch.zeros(self.layout.total, dtype=torch.float32, device=self.device)
            for path, sp in self.layout.leaves.items():
                v = torch.as_tensor(flat[path]).to(self.device, dtype=torch.float32)
                self.layout.view2d(full, path).copy_(v.reshape(sp.rows, sp.cols))
            buf.copy_(self.layout.flat_to_owner_slice(full, self.rank))
            return
        for path, sp in self.layout.leaves.items():
            v2 = torch.as_tensor(flat[path]).to(self.device, dtype=torch.float32).reshape(sp.rows, sp.cols)
            dim = self.shard_dim[path]
            rws, c, ld = self._local_shape2d(path)
            if dim is not None:
                v2 = v2[:, self.sidx * c:(self.sidx + 1) * c] if c != sp.cols else v2[self.sidx * rws:(self.sidx + 1) * rws]
            self._local_view(buf, path).copy_(v2)

    def load_opt_state_global(self, mu: Dict, nu: Dict, count: int) -> None:
        """Inverse of ``opt_state_global``: every rank passes the same global trees and keeps its part."""
        def flatten(tree, prefix=""):
            out = {}
            for k, v in tree.items():
                if isinstance(v, dict):
                    out.update(flatten(v, prefix + k + "/"))
                else:
                    out[prefix + k] = v
            return out
        self._load_moment(self.m, flatten(mu))
        self._load_moment(self.v, flatten(nu))
        self.step_count.fill_(int(count))
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier(group=self.group)

    @staticmethod
    def _nest(flat: Dict[str, object], order: Optional[Sequence[str]] = None) -> Dict:
        out: Dict = {}
        for path in (order or flat.keys()):
            node = out
            *parents, last = path.split("/")
            for p_ in parents:
                node = node.setdefault(p_, {})
            node[last] = flat[path]
        return out

    def opt_state_tree(self) -> Tuple:
        """optax-shaped optimizer state. DP keeps m/v only for the rank's 1/world slice of every
        bucket (owner-update), FSDP for its shards; the tree exposes those flat slices."""
        from ..optim.adamw import EmptyState, ScaleByAdamState
        return (ScaleByAdamState(count=self.step_count, mu={"flat_slice": self.m}, nu={"flat_slice": self.v}),
                EmptyState(), EmptyState())

    # ---------------------------------------------------------------- collectives -------------
    def _make_ctx(self, arena: SymmetricArena, grad: str, param: Optional[str], bf16: Optional[str],
                  signal: str, local_only: bool = False) -> CommCtx:
        ctx = CommCtx()
        ctx.rank, ctx.world = (0, 1) if local_only else (self.rank, self.world)

        def arr(name):
            if local_only:
                a = (C.c_void_p * 1)(arena.peer_ptrs[self.rank] + arena.offsets[name])
            else:
                a = arena.peer_ptr_array(name)
            self._keep.append(a)
            return C.cast(a, C.POINTER(C.c_void_p))

        ctx.peer_grad = arr(grad)
        if param:
            ctx.peer_param = arr(param)
            ctx.mc_param = arena.mc_ptr(param)
        if bf16:
            ctx.peer_bf16 = arr(bf16)
            if param:
                ctx.mc_bf16 = arena.mc_ptr(bf16)
        ctx.peer_signal = arr(signal)
        if local_only:
            ctx.mc_param = ctx.mc_bf16 = None
        else:
            ctx.mc_grad = arena.mc_ptr(grad)
        return ctx

    def _segs_to_device(self, segs: List) -> torch.Tensor:
        raw = b"".join(bytes(s) for s in segs)
        return torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(self.device)

    def _build_segments(self) -> None:
        W, r = self.world, self.rank
        self.bucket_segs: List[Tuple[torch.Tensor, int]] = []
        self._host_segs: List[Seg] = []
        if self.mode == "dp":
            for (b0, b1) in self.layout.bucket_ranges:
                sl = (b1 - b0) // W
                assert sl < 2 ** 31, "bucket slice too large for one fused-update segment: use more buckets"
                seg = Seg(src_off=b0 + r * sl, src_pitch=sl, dst_off=b0 + r * sl, loc_off=b0 // W, rows=1, row_len=sl)
                self.bucket_segs.append((self._segs_to_device([seg]), 1))
                self._host_segs.append(seg)
            return
        per_bucket: Dict[int, List[Seg]] = {}
        for path, sp in self.layout.leaves.items():
            dim = self.shard_dim[path]
            rws, c, ld = self._local_shape2d(path)
            if dim is None:     # replicated leaf: every rank applies the same all-reduced update (fixed summation order)
                seg = Seg(src_off=sp.offset, src_pitch=sp.ld, dst_off=0, loc_off=self.loc_off[path],
                          rows=sp.rows, row_len=sp.ld, flags=_lib.SEG_ORDERED)
            elif c != sp.cols:  # sharded on the last dim: strided rows
                assert ld == c, f"{path}: last-dim shards must be multiples of 8 wide"
                seg = Seg(src_off=sp.offset + self.sidx * c, src_pitch=sp.ld, dst_off=0, loc_off=self.loc_off[path],
                          rows=sp.rows, row_len=c)
            else:               # sharded on rows: contiguous
                seg = Seg(src_off=sp.offset + self.sidx * rws * sp.ld, src_pitch=sp.ld, dst_off=0,
                          loc_off=self.loc_off[path], rows=rws, row_len=sp.ld)
            # the kernel indexes a segment's 16-byte vectors and its rows / columns with 32 bits
            assert seg.rows * max(seg.row_len, seg.src_pitch) < 2 ** 31, \
                f"{path}: segment of {seg.rows} x {seg.row_len} elements is too large for one fused-update segment"
            per_bucket.setdefault(sp.bucket, []).append(seg)
        order = sorted(per_bucket)
        assert order == list(range(len(order))), "bucket ids must be 0..n-1"
        for b in order:
            self.bucket_segs.append((self._segs_to_device(per_bucket[b]), len(per_bucket[b])))
            self._host_segs.extend(per_bucket[b])

    def _sync_all_buckets(self) -> None:
        """All buckets in one launch (used when there is nothing to overlap with: one rank, or overlap off)."""
        if not hasattr(self, "_all_segs"):
            raws = [bytes(t.cpu().numpy().tobytes()) for t, _ in self.bucket_segs]
            self._all_segs = (torch.frombuffer(bytearray(b"".join(raws)), dtype=torch.uint8).to(self.device),
                              sum(n for _, n in self.bucket_segs))
        self._sync_bucket(-1, last=True)

    def _begin_step(self, src: Optional[torch.Tensor], dst: Optional[torch.Tensor]) -> None:
        """First launch of a step: (optional) fp32 -> bf16 cast of the inputs, zeroing of the step
        metrics and of the bias-gradient region, and - with more than one rank - the deferred half of
        the previous update's end barrier: it returns once every peer has published its parameters
        (and therefore has finished reading this rank's gradients), see csrc/comm.cuh."""
        sig = None
        if self.world > 1 and not self.stub_collectives:
            sig = self.arena.peer_ptrs[self.rank] + self.arena.offsets["signal"]
        K.cast_zero(src, dst, self.step_metrics, self._vec_grads, wait_signal=sig, peer_signals=self.ctx.peer_signal,
                    rank=self.rank, world=self.world)

    def _end_step(self, next_rng) -> None:
        """state.rng <- rebase(next_rng) (the salt folds into the device-resident key in place) and
        step counter += 1. Normally this rides on the step's last fused update (``_fold_end``); if no
        update consumed it, it is one small launch."""
        if self._end_folded:
            self._end_folded = False
        else:
            K.rng_split_bump_(self.rng_base, _rng.salt_word(next_rng), self.step_count)
        self._end_salt = None
        self.release_gathered()

    _end_salt = None
    _end_folded = False

    def _fold_end(self, next_rng) -> None:
        """Call before the step's last ``_sync_bucket(..., last=True)``: that launch then also advances
        the RNG key and the step counter (the block that finishes last does it)."""
        if not hasattr(self, "_done_word"):
            self._done_word = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._end_salt = _rng.salt_word(next_rng)

    def _sync_bucket(self, b: int, last: bool) -> None:
        """Fused reduce + mean + AdamW + publish of one gradient bucket (one launch)."""
        segs, nseg = self._all_segs if b < 0 else self.bucket_segs[b]
        hp = AdamWHParams(self.hp["lr"], self.hp["beta1"], self.hp["beta2"], self.hp["eps"],
                          self.hp["weight_decay"], 1.0 / (self.num_minibatches * self.world))
        nmet = len(self.metric_names)
        dp = self.mode == "dp"
        end = None
        if last and self._end_salt is not None:
            end = _lib.StepEnd(self.rng_base.data_ptr(), self._end_salt, self.step_count.data_ptr(), self._done_word.data_ptr())
            self._end_folded = True
        u = _lib.FusedUpdateArgs()
        u.ctx = C.pointer(self.ctx_stub if self.stub_collectives else self.ctx)
        u.segs_dev, u.nseg, u.broadcast = segs.data_ptr(), nseg, int(dp)
        u.publish_fp32 = int(self.publish_fp32)
        u.nmetrics = nmet if last else 0
        u.blocks = self.comm_blocks
        u.p_local = None if dp else self.p_loc.data_ptr()
        u.bf16_local = None if dp else self.pb.data_ptr()
        u.m, u.v = self.m.data_ptr(), self.v.data_ptr()
        u.hp = C.pointer(hp)
        u.step_count = self.step_count.data_ptr()     # bumped by the end record / _end_step
        u.step_metrics = self.step_metrics.data_ptr() if last else None
        u.metrics_acc = self.metrics_acc.data_ptr() if last else None
        u.end = C.pointer(end) if end is not None else None
        _lib.check(_lib.lib().b200_fused_update(C.byref(u), torch.cuda.current_stream().cuda_stream))
        if dp and not self.publish_fp32:
            self._master_stale = True

    def _sync_bucket_async(self, b: int) -> None:
        """Launches bucket ``b``'s fused update on the side stream, after everything already
        queued on the current stream (its gradients), so it runs under the rest of the backward.
        The fused launches share the signal words, hence they are serialised on the side stream."""
        main = torch.cuda.current_stream()
        self.side_stream.wait_stream(main)
        with torch.cuda.stream(self.side_stream):
            self._sync_bucket(b, last=False)

    # ---------------------------------------------------------------- self-check / measurement aids
    def set_stub_collectives(self, on: bool) -> None:
        """Switches the "collectives stubbed" variant of the step on or off (bench.py's exposed-comm
        measurement): every cross-GPU access is replaced by the same access to local memory and the
        flag exchanges are skipped, so the kernels do the same arithmetic and local traffic without
        NVLink and without waiting for peers. The parameters it produces are NOT a training result.
        The step graph is re-captured on the next call."""
        on = bool(on) and self.world > 1
        if on == self.stub_collectives:
            return
        self.stub_collectives = on
        if on and not hasattr(self, "ctx_stub"):
            self.ctx_stub = self._make_ctx(self.arena, "grad", "param" if self.mode == "dp" else None, "bf16", "signal",
                                           local_only=True)
        self.graph = None
        self._warmed = False
        self.release_gathered()
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier(group=self.group)

    def _peer_base(self, g: int) -> int:
        """Base address of rank ``g``'s arena as this rank sees it (its own when collectives are stubbed)."""
        return self.arena.peer_ptrs[self.rank if self.stub_collectives else g]

    def selfcheck(self, run_one_step) -> Dict[str, object]:
        """One real step of the fused update checked against an independent path: the gradients the
        step left in ``g32`` are all-reduced with NCCL, a plain fp32 PyTorch AdamW is applied to a
        snapshot of this rank's master parameters / moments taken before the step, and the result is
        compared with what the fused kernel wrote; the bf16 working copies (DP) must be bit-identical
        on every rank. ``run_one_step`` is a callable that performs exactly one training step.
        Collective: every rank calls it. Returns {"ok", "max_abs_err", "replicas_identical"}."""
        dp = self.mode == "dp"
        torch.cuda.synchronize()
        p_buf = self.p32 if dp else self.p_loc
        p0, m0, v0 = p_buf.clone(), self.m.clone(), self.v.clone()
        t = int(self.step_count.item()) + 1
        run_one_step()
        torch.cuda.synchronize()
        g = self.g32.clone()
        if self.world > 1:
            dist.all_reduce(g, group=self.group)
        g.mul_(1.0 / (self.num_minibatches * self.world))
        hp = self.hp
        worst = 0.0
        for sg in self._host_segs:
            rows, rl = int(sg.rows), int(sg.row_len)
            src = torch.as_strided(g, (rows, rl), (int(sg.src_pitch), 1), int(sg.src_off)).reshape(-1)
            loc = slice(int(sg.loc_off), int(sg.loc_off) + rows * rl)
            pos = slice(int(sg.dst_off), int(sg.dst_off) + rows * rl) if dp else loc
            m = hp["beta1"] * m0[loc] + (1 - hp["beta1"]) * src
            v = hp["beta2"] * v0[loc] + (1 - hp["beta2"]) * src * src
            mhat, vhat = m / (1 - hp["beta1"] ** t), v / (1 - hp["beta2"] ** t)
            ref = p0[pos] - hp["lr"] * (mhat / (vhat.sqrt() + hp["eps"]) + hp["weight_decay"] * p0[pos])
            worst = max(worst, (p_buf[pos] - ref).abs().max().item(), (self.m[loc] - m).abs().max().item())
        same = True
        if dp and self.world > 1:
            ref_b = self.pb.clone()
            src0 = dist.get_global_rank(self.group, 0) if self.group is not dist.group.WORLD else 0
            dist.broadcast(ref_b, src0, group=self.group)
            flag = torch.tensor([int(torch.equal(ref_b, self.pb))], device=self.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
            same = bool(flag.item())
        w = torch.tensor([worst], device=self.device, dtype=torch.float64)
        if self.world > 1:
            dist.all_reduce(w, op=dist.ReduceOp.MAX, group=self.group)
        worst = float(w.item())
        return {"ok": bool(worst < 1e-5 and same), "max_abs_err": worst, "replicas_identical": same,
                "against": "NCCL all_reduce of the step's gradients + fp32 PyTorch AdamW"}

    # ---------------------------------------------------------------- weights for GEMMs -------
    def grad2d(self, path: str) -> torch.Tensor:
        return self.layout.view2d(self.g32, path)

    def grad1d(self, path: str) -> torch.Tensor:
        return self.layout.view(self.g32, path)

    def vec(self, path: str) -> torch.Tensor:
        """bf16 working copy of a 1-D leaf (bias / norm scale), full length."""
        if self.mode == "dp":
            return self.layout.view(self.pb, path)
        if self.shard_dim[path] is None:
            return self._local_view(self.pb, path).reshape(-1)
        return self.bias_full[path]

    def _weight(self, path: str):
        sp = self.layout.leaves[path]
        if self.mode == "dp":
            return self.layout.view2d(self.pb, path), {}, None
        if self.shard_dim[path] is None or self.S == 1:
            return self._local_view(self.pb, path), {}, None
        if self.fsdp_weights == "gather":
            return self._gathered(path), {}, None
        r, c, ld = self._local_shape2d(path)
        srcs = [self._peer_base(g) + self.arena.offsets["bf16"] + self.loc_off[path] * 2
                for g in self.shard_peer_ranks]
        last_dim = c != sp.cols
        return srcs, dict(b_shape=(sp.rows, sp.cols), ldb=ld), (last_dim, c if last_dim else r)

    def gemm_w(self, x: torch.Tensor, path: str, out: torch.Tensor, **kw) -> None:
        """``out = epilogue(x @ W)`` with W the (in, out) kernel at ``path``. In FSDP mode W is read
        from the peers' shards by the GEMM's TMA producer (all-gather fused into the GEMM;
        reference: gather_params then Dense, param_sharding.py:146-175)."""
        b, extra, shard = self._weight(path)
        if shard is not None:
            last_dim, shard_len = shard
            extra.update(shard_mode=2 if last_dim else 1, shard_len=shard_len, rot=self.sidx)
        K.gemm(x, b, out, b_mn=True, **extra, **kw)

    def gemm_wT(self, dy: torch.Tensor, path: str, out: torch.Tensor, **kw) -> None:
        """``out = epilogue(dy @ W^T)`` (dgrad): the (in, out) kernel is the K-major B operand."""
        b, extra, shard = self._weight(path)
        if shard is not None:
            last_dim, shard_len = shard
            extra.update(shard_mode=1 if last_dim else 2, shard_len=shard_len, rot=self.sidx)
        K.gemm(dy, b, out, **extra, **kw)

    # ---------------------------------------------------------------- step driver -------------
    def _load_batch(self, batch: Batch) -> None:  # pragma: no cover - interface
        raise NotImplementedError

    def _step_body(self) -> None:  # pragma: no cover - interface
        raise NotImplementedError

    def _rows_per_step(self) -> int:  # pragma: no cover - interface
        raise NotImplementedError

    # Programmatic dependent launch between the step's kernels: pays for chains of short kernels
    # (the MLP step is launch-latency bound), costs ~2% on steps made of long persistent kernels
    # (measured on GPT-2 small, 16k tokens per step). Engines set this from their step size.
    pdl = True

    def _launch_body(self) -> None:
        _lib.set_pdl(1 if self.pdl else 0)
        try:
            self._step_body()
        finally:
            _lib.set_pdl(-1)

    def _run_step(self) -> None:
        """Call 1 runs the step eagerly (lazy initialisation inside CUDA/NCCL happens here), call 2
        captures it into a CUDA graph and replays it, later calls only replay. Every call performs
        exactly one real training step."""
        if not self.use_graph:
            self._master_stale = True
            before = _lib.launch_count()
            self._launch_body()
            self.launches_per_step = _lib.launch_count() - before
            return
        self._master_stale = True      # (replays do not pass through _sync_bucket)
        if self.graph is None:
            if not getattr(self, "_warmed", False):
                s = torch.cuda.Stream()
                s.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(s):
                    before = _lib.launch_count()
                    self._launch_body()
                    self.launches_per_step = _lib.launch_count() - before
                torch.cuda.current_stream().wait_stream(s)
                self._warmed = True
                return
            torch.cuda.synchronize()
            if self.world > 1:
                dist.barrier(group=self.group)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=torch.cuda.Stream()):
                self._launch_body()
            self.graph = g
        self.graph.replay()

    def train_step(self, state: TrainState, metrics: Optional[Metrics], batch: Batch) -> Tuple[TrainState, Metrics]:
        """``(state, metrics, batch) -> (state, metrics)`` like the reference's jitted steps. The
        state's buffers are updated in place; the returned state replaces the argument."""
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
        return dataclasses.replace(state, step=state.step + 1), out

    def make_state(self, model, rng: _rng.Key, params_tree: Optional[Dict] = None) -> TrainState:
        from ..optim import adamw as _adamw
        self.rng_base = rng.base.to(self.device).clone()
        if rng.salt != 0:
            self.rng_base = _rng.rebase(_rng.Key(self.rng_base, rng.salt)).base
        key = _rng.Key(self.rng_base, 0)
        return TrainState(step=0, apply_fn=model.apply, params=params_tree if params_tree is not None else self.params_tree(gather=False),
                          tx=_adamw(self.hp["lr"]), opt_state=self.opt_state_tree(), rng=key, engine=self)

    def params_tree(self, gather: bool = True) -> Dict:
        return self._nest(self.flat_params(gather=gather))

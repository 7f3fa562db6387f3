"""Fused engine for the reference's MLP classifier in data-parallel and FSDP modes.

Per minibatch (reference: util.py:58-75 around data_paral.py:171-189) the schedule is eight
launches, all ours:

    GEMM  h = dropout(silu(x W1 + b1))        bias+SiLU+dropout in the epilogue, pre-act saved
    GEMM  logits = h W2 + b2                  fp32 out
    xent  loss sum, #correct, dlogits         one pass
    GEMM  dW2 (+)= h^T dlogits                accumulates straight into the flat gradient buffer
    colsum db2 (+)= sum dlogits
    GEMM  dpre = (dlogits W2^T) * mask/keep * silu'(pre)     backward epilogue
    GEMM  dW1 (+)= x^T dpre
    colsum db1 (+)= sum dpre

then one fused collective+optimizer launch per gradient bucket:

    DP   (data_paral.py:210-217)        NVLS all-reduce-mean + AdamW on the rank's slice +
                                        multicast publish of fp32/bf16 params; bucket 0
                                        (output layer) runs on a side stream under the rest of
                                        the last minibatch's backward.
    FSDP (param_sharding.py:134-140,353) the weight GEMMs read the peers' bf16 shards in place
                                        through TMA (all-gather fused into the GEMM); the
                                        reduce-scatter-mean + sharded AdamW is one launch.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from .. import rng as _rng
from ..mesh import Mesh
from ..models.mlp import Classifier
from ..ops import kernels as K
from ..ops._lib import Seg
from ..parallel.fsdp import choose_shard_dim
from ..state import Batch, TrainState
from ..tree import Partitioned
from ..optim import adamw as _adamw
from ..optim.adamw import EmptyState, ScaleByAdamState
from .base import FlatLayout, FusedEngine, SymmetricArena, _round_up

_LEAVES = ("input_dense/kernel", "input_dense/bias", "output_dense/kernel", "output_dense/bias")


class MLPEngine(FusedEngine):
    def __init__(self, config, mesh: Mesh, *, mode: str = "dp", batch_per_rank: int, lr: float,
                 num_minibatches: int = 4, min_weight_size: int = 2 ** 18, seed_key: Optional[_rng.Key] = None,
                 init_params: Optional[Dict] = None, input_size: int = 784, overlap: bool = True,
                 fuse_minibatches: bool = True, **kw):
        super().__init__(mesh, lr=lr, num_minibatches=num_minibatches,
                         data_axis_name=config.get("data_axis_name", "data"), **kw)
        assert mode in ("dp", "fsdp")
        self.mode, self.cfg = mode, config
        self.B = batch_per_rank
        assert self.B % num_minibatches == 0
        self.mb = self.B // num_minibatches
        self.In, self.H, self.Cn = input_size, config.hidden_size, config.num_classes
        self.rate = float(config.dropout_rate)
        self.overlap = overlap and self.world > 1
        # Gradient accumulation folded into the GEMMs: all minibatches go through one set of
        # launches, the wgrad K loop does the summation and the dropout epilogue switches RNG
        # stream per minibatch row block, so results equal the reference's loop (util.py:58-77).
        self.fuse_minibatches = fuse_minibatches and num_minibatches <= 16
        self.min_weight_size = min_weight_size
        W = self.world
        # bucket 0 = output layer (gradients complete first in backward), bucket 1 = input layer
        self.layout = FlatLayout([
            ("output_dense/kernel", (self.H, self.Cn), 0),
            ("output_dense/bias", (self.Cn,), 0),
            ("input_dense/kernel", (self.In, self.H), 1),
            ("input_dense/bias", (self.H,), 1),
        ], W)
        n = self.layout.total
        self._alloc_common()
        d = self.device
        sizes = {"grad": n * 4, "signal": 8192 * 4}
        if mode == "dp":
            sizes.update({"param": n * 4, "bf16": n * 2})
        else:
            self._plan_fsdp_shards()
            sizes.update({"bf16": self.local_total * 2})   # every rank's bf16 slices, same local offsets
        self.arena = SymmetricArena(sizes, mesh, d)
        self.g32 = self.arena.tensor("grad", torch.float32, n)
        if mode == "dp":
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
        self.ctx = self._make_ctx(self.arena, "grad", "param" if mode == "dp" else None, "bf16", "signal")
        self._build_segments()
        # activations (static, reused every minibatch)
        mb = self.B if self.fuse_minibatches else self.mb
        pad8 = lambda c: _round_up(c, 8)
        self.x_in = torch.zeros(self.B, self.In, dtype=torch.float32, device=d)
        self.labels = torch.zeros(self.B, dtype=torch.int64, device=d)
        self.x_bf = torch.zeros(self.B, pad8(self.In), dtype=torch.bfloat16, device=d)
        self.pre = torch.zeros(mb, self.H, dtype=torch.bfloat16, device=d)
        self.h = torch.zeros(mb, self.H, dtype=torch.bfloat16, device=d)
        self.logits = torch.zeros(mb, pad8(self.Cn), dtype=torch.float32, device=d)
        self.dlogits = torch.zeros(mb, pad8(self.Cn), dtype=torch.bfloat16, device=d)
        self.dpre = torch.zeros(mb, self.H, dtype=torch.bfloat16, device=d)
        # parameters
        self.rng_base = None
        if init_params is not None:
            self.load_params(init_params)

    # ------------------------------------------------------------------ parameters / state ---
    def _plan_fsdp_shards(self) -> None:
        """Shard decisions follow ``shard_params`` (param_sharding.py:89-113). Sharded leaves keep
        fp32 master / m / v / bf16 for the local slice only; the bf16 slices live in the symmetric
        arena at the same local offset on every rank, where the peers' GEMMs read them through
        TMA. Replicated leaves (too small or indivisible) are kept whole on every rank and every
        rank applies the same (all-reduced) update to them."""
        W = self.world
        self.shard_dim: Dict[str, Optional[int]] = {}
        loc = 0
        self.loc_off: Dict[str, int] = {}
        for path, sp in self.layout.leaves.items():
            numel = 1
            for s in sp.shape:
                numel *= s
            dim = None
            if numel > self.min_weight_size and W > 1:
                dim = choose_shard_dim(sp.shape, (None,) * len(sp.shape), W)
            elif W == 1 and numel > self.min_weight_size:
                dim = choose_shard_dim(sp.shape, (None,) * len(sp.shape), 1)
            self.shard_dim[path] = dim
            self.loc_off[path] = loc
            loc = _round_up(loc + self._local_numel(path), 64)
        self.local_total = _round_up(loc, 64)

    def _setup_bias_gather(self) -> None:
        """Sharded 1-D leaves (biases) are re-assembled once per step by peer loads."""
        import ctypes as C
        from ..ops._lib import GatherSeg
        self.bias_full: Dict[str, torch.Tensor] = {}
        segs, off = [], 0
        for path, sp in self.layout.leaves.items():
            if len(sp.shape) == 1 and self.shard_dim[path] is not None:
                per = sp.cols // self.world
                segs.append(GatherSeg(self.loc_off[path] * 2, off, per * 2))
                self.bias_full[path] = (off // 2, sp.cols)
                off = _round_up(off + sp.cols * 2, 64)
        self._bias_nseg = len(segs)
        if segs:
            self._bias_buf = torch.zeros(off // 2, dtype=torch.bfloat16, device=self.device)
            self._bias_segs = self._segs_to_device(segs)
            self._bias_peers = self.arena.peer_ptr_array("bf16")
            self.bias_full = {k: self._bias_buf[o: o + n] for k, (o, n) in self.bias_full.items()}

    def _gather_biases(self) -> None:
        if not self._bias_nseg:
            return
        import ctypes as C
        from ..ops import _lib
        _lib.check(_lib.lib().b200_peer_gather(C.cast(self._bias_peers, C.POINTER(C.c_void_p)), self.world,
                                                self._bias_segs.data_ptr(), self._bias_nseg,
                                                self._bias_buf.data_ptr(), 1,
                                                torch.cuda.current_stream().cuda_stream))

    def _local_shape2d(self, path: str) -> Tuple[int, int, int]:
        """(rows, cols, ld) of this rank's slice of a leaf in padded 2-D form."""
        sp = self.layout.leaves[path]
        dim = self.shard_dim[path]
        W = self.world
        if dim is None:
            return sp.rows, sp.cols, sp.ld
        if len(sp.shape) == 1 or dim == len(sp.shape) - 1:
            cols = sp.cols // W
            return sp.rows, cols, _round_up(cols, 8)
        return sp.rows // W, sp.cols, sp.ld

    def _local_numel(self, path: str) -> int:
        r, _, ld = self._local_shape2d(path)
        return r * ld

    def _local_view(self, buf: torch.Tensor, path: str, base: Optional[int] = None) -> torch.Tensor:
        r, c, ld = self._local_shape2d(path)
        o = self.loc_off[path] if base is None else base
        return buf[o: o + r * ld].view(r, ld)[:, :c]

    def load_params(self, params: Dict) -> None:
        """Loads a *global* (unsharded) reference-layout parameter tree; every rank passes the same
        values (the reference initialises the full parameter on every device from one rng and
        keeps its slice, param_sharding.py:227-241)."""
        flat = {f"{a}/{b}": v for a, sub in params.items() for b, v in sub.items()}
        if self.mode == "dp":
            for path in self.layout.leaves:
                v = flat[path]
                v = v.value if isinstance(v, Partitioned) else v
                self.layout.view2d(self.p32, path).copy_(v.reshape(self.layout.leaves[path].rows, -1))
            K.cast_f32_bf16(self.p32, self.pb)
        else:
            for path, sp in self.layout.leaves.items():
                v = flat[path]
                v = (v.value if isinstance(v, Partitioned) else v).to(self.device)
                v2 = v.reshape(sp.rows, sp.cols)
                dim = self.shard_dim[path]
                r, c, ld = self._local_shape2d(path)
                if dim is not None:
                    if c != sp.cols:
                        v2 = v2[:, self.rank * c:(self.rank + 1) * c]
                    else:
                        v2 = v2[self.rank * r:(self.rank + 1) * r]
                self._local_view(self.p_loc, path).copy_(v2)
                self._local_view(self.pb, path).copy_(v2)
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()

    def params_tree(self, gather: bool = True) -> Dict:
        """Reference-layout parameter tree. DP: views of the replicated fp32 master. FSDP:
        ``Partitioned`` boxes of the local slices (``gather=False``) or full arrays."""
        if self.mode == "dp":
            return self.layout.tree(self.p32)
        out: Dict = {}
        for path, sp in self.layout.leaves.items():
            a, b = path.split("/")
            loc = self._local_view(self.p_loc, path)
            dim = self.shard_dim[path]
            loc = loc.reshape(-1) if len(sp.shape) == 1 else loc
            if dim is None:
                leaf = loc
            elif gather:
                from ..comm import collectives as C
                leaf = C.all_gather(loc.contiguous(), self.data_axis_name, axis=dim, mesh=self.mesh)
            else:
                names = tuple(self.data_axis_name if i == dim else None for i in range(len(sp.shape)))
                leaf = Partitioned(loc, names)
            out.setdefault(a, {})[b] = leaf
        # present leaves in the reference's order
        return {k: {kk: out[k][kk] for kk in ("kernel", "bias")} for k in ("input_dense", "output_dense")}

    def opt_state_tree(self) -> Tuple:
        """optax-shaped view of the optimizer state. DP keeps m/v only for this rank's 1/world
        slice of each bucket (owner-update), so the tree exposes flat slices."""
        return (ScaleByAdamState(count=self.step_count, mu={"flat_slice": self.m}, nu={"flat_slice": self.v}),
                EmptyState(), EmptyState())

    def make_state(self, model: Classifier, rng: _rng.Key) -> TrainState:
        self.rng_base = rng.base.to(self.device).clone()
        self._rng_salt0 = rng.salt
        if rng.salt != 0:
            self.rng_base = _rng.rebase(_rng.Key(self.rng_base, rng.salt)).base
        key = _rng.Key(self.rng_base, 0)
        return TrainState(step=0, apply_fn=model.apply, params=self.params_tree(gather=False),
                          tx=_adamw(self.hp["lr"]), opt_state=self.opt_state_tree(), rng=key, engine=self)

    # ------------------------------------------------------------------ collectives ----------
    def _build_segments(self) -> None:
        W, r = self.world, self.rank
        self.bucket_segs: List[Tuple[torch.Tensor, int]] = []
        if self.mode == "dp":
            for (b0, b1) in self.layout.bucket_ranges:
                sl = (b1 - b0) // W
                seg = Seg(src_off=b0 + r * sl, src_pitch=sl, dst_off=b0 + r * sl, loc_off=b0 // W, rows=1, row_len=sl)
                self.bucket_segs.append((self._segs_to_device([seg]), 1))
            return
        # FSDP: sharded leaves -> reduce my slice region of every peer's full gradient (local
        # update); replicated leaves -> DP-style but updated redundantly by every rank (the
        # reference pmeans them, param_sharding.py:316-318, then every rank applies the update)
        per_bucket: Dict[int, List[Seg]] = {}
        for path, sp in self.layout.leaves.items():
            dim = self.shard_dim[path]
            rws, c, ld = self._local_shape2d(path)
            if dim is None:
                seg = Seg(src_off=sp.offset, src_pitch=sp.ld, dst_off=0, loc_off=self.loc_off[path],
                          rows=sp.rows, row_len=sp.ld)
            elif c != sp.cols:   # sharded on the last dim: strided rows
                seg = Seg(src_off=sp.offset + r * c, src_pitch=sp.ld, dst_off=0, loc_off=self.loc_off[path],
                          rows=sp.rows, row_len=c)
                assert ld == c, "last-dim shards must be multiples of 8 wide"
            else:                # sharded on rows: contiguous
                seg = Seg(src_off=sp.offset + r * rws * sp.ld, src_pitch=sp.ld, dst_off=0,
                          loc_off=self.loc_off[path], rows=rws, row_len=sp.ld)
            per_bucket.setdefault(sp.bucket, []).append(seg)
        for b in sorted(per_bucket):
            self.bucket_segs.append((self._segs_to_device(per_bucket[b]), len(per_bucket[b])))

    def _sync_bucket(self, b: int, last: bool) -> None:
        segs, nseg = self.bucket_segs[b]
        scale = 1.0 / (self.num_minibatches * self.world)
        if self.mode == "dp":
            self._fused_update(self.ctx, segs, nseg, broadcast=True, p_local=None, bf16_local=None,
                               m=self.m, v=self.v, grad_scale=scale, bump=last, with_metrics=last)
        else:
            self._fused_update(self.ctx, segs, nseg, broadcast=False, p_local=self.p_loc, bf16_local=self.pb,
                               m=self.m, v=self.v, grad_scale=scale, bump=last, with_metrics=last)

    # ------------------------------------------------------------------ weights for GEMMs ----
    def _weight(self, path: str):
        """What the GEMM needs to read a (possibly sharded) weight: kwargs for ``K.gemm``'s B."""
        sp = self.layout.leaves[path]
        if self.mode == "dp":
            return dict(b=self.layout.view2d(self.pb, path)), None
        if self.shard_dim[path] is None:
            return dict(b=self._local_view(self.pb, path)), None
        r, c, ld = self._local_shape2d(path)
        srcs = [p + self.arena.offsets["bf16"] + self.loc_off[path] * 2 for p in self.arena.peer_ptrs]
        return dict(b=srcs, b_shape=(sp.rows, sp.cols), ldb=ld), (c != sp.cols, r if c == sp.cols else c)

    # ------------------------------------------------------------------ the step --------------
    def _load_batch(self, batch: Batch) -> None:
        self.x_in.copy_(batch.inputs, non_blocking=True)
        self.labels.copy_(batch.labels, non_blocking=True)

    def _rows_per_step(self) -> int:
        return self.B * self.world

    def _step_body(self) -> None:
        nmb, mb = self.num_minibatches, self.mb
        key = _rng.Key(self.rng_base, 0)
        next_rng, step_rng = _rng.split(key, 2)
        mb_keys = _rng.split(step_rng, nmb)
        axis_idx = self.mesh.axis_index(self.data_axis_name)
        salts = [_rng.salt_word(_rng.fold_in(mb_keys[i], axis_idx)) for i in range(nmb)]
        K.scale_(self.step_metrics, 0.0)
        xb = self.x_bf[:, : self.In]
        if self.x_bf.shape[1] == self.In:
            K.cast_f32_bf16(self.x_in, self.x_bf)
        else:
            xb.copy_(self.x_in)
        if self.mode == "fsdp":
            self._gather_biases()
        if self.fuse_minibatches:
            self._fwd_bwd(xb, self.labels, salts, mb, accumulate=False, last=True)
        else:
            for i in range(nmb):
                self._fwd_bwd(xb[i * mb:(i + 1) * mb], self.labels[i * mb:(i + 1) * mb], [salts[i]], mb,
                              accumulate=i > 0, last=i == nmb - 1)
        if self.overlap:
            torch.cuda.current_stream().wait_stream(self.side_stream)   # the fused launches share signal words
        else:
            self._sync_bucket(0, last=False)
        self._sync_bucket(1, last=True)
        # state.rng <- rebase(next_rng): the salt folds into the device-resident key in place
        K.rng_split_(self.rng_base, None, _rng.salt_word(next_rng), 0)

    def _fwd_bwd(self, xs, lab, salts, rows_per_salt: int, *, accumulate: bool, last: bool) -> None:
        """Forward + backward of one row block (a minibatch, or all minibatches at once)."""
        lay = self.layout
        g = lambda p: lay.view2d(self.g32, p)
        logits = self.logits[:, : self.Cn]
        dlogits = self.dlogits[:, : self.Cn]
        drop = dict(drop_key=self.rng_base, drop_rate=self.rate, drop_salts=salts, drop_rows_per_salt=rows_per_salt)
        main = torch.cuda.current_stream()
        if self.mode == "dp":
            W1, W2 = lay.view2d(self.pb, "input_dense/kernel"), lay.view2d(self.pb, "output_dense/kernel")
            b1, b2 = lay.view(self.pb, "input_dense/bias"), lay.view(self.pb, "output_dense/bias")
            K.gemm(xs, W1, self.h, b_mn=True, bias=b1, act="silu", pre=self.pre, **drop)
            K.gemm(self.h, W2, logits, b_mn=True, bias=b2)
        else:
            self._fsdp_gemm_fwd(xs, "input_dense", self.h, act="silu", pre=self.pre, drop=drop)
            self._fsdp_gemm_fwd(self.h, "output_dense", logits)
        K.softmax_xent(logits, lab, grad_scale=1.0 / self.mb, dlogits=dlogits, metrics=self.step_metrics)
        K.gemm(self.h, dlogits, g("output_dense/kernel"), a_mn=True, b_mn=True, accumulate=accumulate)
        K.colsum(dlogits, lay.view(self.g32, "output_dense/bias"), accumulate=accumulate)
        if last and self.overlap:
            # bucket 0 (output layer) is complete: reduce + update it under the rest of the backward
            self.side_stream.wait_stream(main)
            with torch.cuda.stream(self.side_stream):
                self._sync_bucket(0, last=False)
        if self.mode == "dp":
            K.gemm(dlogits, W2, self.dpre, bwd=True, aux=self.pre, act="silu", **drop)
        else:
            self._fsdp_gemm_dgrad(dlogits, "output_dense", self.dpre, aux=self.pre, act="silu", drop=drop)
        K.gemm(xs, self.dpre, g("input_dense/kernel"), a_mn=True, b_mn=True, accumulate=accumulate)
        K.colsum(self.dpre, lay.view(self.g32, "input_dense/bias"), accumulate=accumulate)

    # FSDP GEMMs: B comes from the peers' shards (all-gather fused into the GEMM)
    def _fsdp_gemm_fwd(self, x, layer: str, out, act=None, pre=None, drop=None):
        kw, shard = self._weight(f"{layer}/kernel")
        bias = self._bias(layer)
        extra = {}
        if shard is not None:
            last_dim, shard_len = shard
            extra = dict(shard_mode=2 if last_dim else 1, shard_len=shard_len, rot=self.rank)
        b = kw.pop("b")
        K.gemm(x, b, out, b_mn=True, bias=bias, act=act, pre=pre, **(drop or {}), **kw, **extra)

    def _fsdp_gemm_dgrad(self, dy, layer: str, out, aux=None, act=None, drop=None):
        kw, shard = self._weight(f"{layer}/kernel")
        extra = {}
        if shard is not None:
            last_dim, shard_len = shard
            # dx = dy W^T: the (in,out) kernel is the K-major B operand [N'=in][K'=out]
            extra = dict(shard_mode=1 if last_dim else 2, shard_len=shard_len, rot=self.rank)
        b = kw.pop("b")
        K.gemm(dy, b, out, bwd=True, aux=aux, act=act, **(drop or {}), **kw, **extra)

    def _bias(self, layer: str) -> torch.Tensor:
        """Biases are tiny: sharded ones are gathered once per step into a local bf16 vector by the
        generic collective (an 1 KB all-gather); replicated ones are read in place."""
        path = f"{layer}/bias"
        if self.shard_dim[path] is None:
            return self._local_view(self.pb, path).reshape(-1)
        return self.bias_full[path]

"""Fused engine for the reference's MLP classifier in data-parallel and FSDP modes.

One step (all minibatches folded into the row dimension, reference: util.py:58-75 around
data_paral.py:171-189) is seven launches, all ours, replayed as one CUDA graph:

    cast  x -> bf16, zero the step metrics and the bias-gradient vectors
    GEMM  h = dropout(silu(x W1 + b1))        bias+SiLU+dropout in the epilogue, pre-act saved
    GEMM  dlogits = xent'(h W2 + b2)          softmax cross-entropy in the epilogue (N <= 32): loss sum,
                                              #correct, dlogits and db2 leave the kernel, logits do not
    GEMM  dW2 = h^T dlogits                   parallel graph branch (side stream)
    GEMM  dpre = (dlogits W2^T) * mask/keep * silu'(pre)     backward epilogue, db1 as column sums
    GEMM  dW1 = x^T dpre
    fused collective + optimizer + end-of-step bookkeeping (RNG key, step counter)

    DP   (data_paral.py:210-217)        NVLS all-reduce-mean + AdamW on the rank's slice +
                                        multicast publish of fp32/bf16 params. Models above 8 MB
                                        use one launch per bucket, the output layer's on a side
                                        stream under the rest of the backward.
    FSDP (param_sharding.py:134-140,353) the weight GEMMs read the peers' bf16 shards in place
                                        through TMA (all-gather fused into the GEMM); the
                                        reduce-scatter-mean + sharded AdamW is one launch.

With more than 32 classes the head and the cross-entropy are separate launches; without
``fuse_minibatches`` the GEMM block runs once per minibatch and accumulates.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from .. import rng as _rng
from ..mesh import Mesh
from ..ops import kernels as K
from ..state import Batch
from .base import FlatLayout, FusedEngine, _round_up

_ORDER = ("input_dense/kernel", "input_dense/bias", "output_dense/kernel", "output_dense/bias")


class MLPEngine(FusedEngine):
    def __init__(self, config, mesh: Mesh, *, mode: str = "dp", batch_per_rank: int, lr: float,
                 num_minibatches: int = 4, min_weight_size: int = 2 ** 18, init_params: Optional[Dict] = None,
                 input_size: int = 784, fuse_minibatches: bool = True, fuse_xent: bool = True, **kw):
        # every pass of the fused collective's loop costs one NVLink round trip (~5 us measured): enough
        # blocks to cover the whole model in a single pass beats saving barrier participants
        kw.setdefault("comm_blocks", 128)
        if kw.get("fsdp_weights", "auto") == "auto":   # few row tiles per GEMM: read the peers' shards from inside the GEMM
            kw["fsdp_weights"] = "fused" if batch_per_rank <= 1024 else "gather"
        super().__init__(mesh, mode=mode, lr=lr, num_minibatches=num_minibatches,
                         data_axis_name=config.get("data_axis_name", "data"), min_weight_size=min_weight_size, **kw)
        self.cfg = config
        self.B = batch_per_rank
        assert self.B % num_minibatches == 0
        self.mb = self.B // num_minibatches
        self.In, self.H, self.Cn = input_size, config.hidden_size, config.num_classes
        self.rate = float(config.dropout_rate)
        # Gradient accumulation folded into the GEMMs: all minibatches go through one set of
        # launches, the wgrad K loop does the summation and the dropout epilogue switches RNG
        # stream per minibatch row block, so results equal the reference's loop (util.py:58-77).
        self.fuse_minibatches = fuse_minibatches and num_minibatches <= 16
        # bucket 0 = output layer's matrix (its gradient completes first in backward), bucket 1 = input
        # layer's matrix + both bias vectors (contiguous: cleared with one launch, filled by atomics
        # from the xent kernel and the dgrad epilogue instead of separate column-sum passes)
        self._setup_buffers(FlatLayout([
            ("output_dense/kernel", (self.H, self.Cn), 0),
            ("input_dense/kernel", (self.In, self.H), 1),
            ("input_dense/bias", (self.H,), 1),
            ("output_dense/bias", (self.Cn,), 1),
        ], self.world))
        # a sub-megabyte model: one fused reduce+update launch (two cross-GPU barriers) beats two
        # overlapped launches (four barriers); bucket-level overlap pays off from a few MB per bucket
        self.overlap = self.overlap and self.layout.total * 4 > (8 << 20)
        self.fork_wgrad = (not self.overlap) and self.fuse_minibatches
        self.fuse_xent = fuse_xent
        b1, b2 = self.layout.leaves["input_dense/bias"], self.layout.leaves["output_dense/bias"]
        self._vec_grads = self.g32[b1.offset: b2.offset + b2.ld]
        d = self.device
        rows = self.B if self.fuse_minibatches else self.mb
        pad8 = lambda c: _round_up(c, 8)
        self.x_in = torch.zeros(self.B, self.In, dtype=torch.float32, device=d)
        self.labels = torch.zeros(self.B, dtype=torch.int64, device=d)
        self.x_bf = torch.zeros(self.B, pad8(self.In), dtype=torch.bfloat16, device=d)
        self.pre = torch.zeros(rows, self.H, dtype=torch.bfloat16, device=d)
        self.h = torch.zeros(rows, self.H, dtype=torch.bfloat16, device=d)
        self.logits = torch.zeros(rows, pad8(self.Cn), dtype=torch.float32, device=d)
        self.dlogits = torch.zeros(rows, pad8(self.Cn), dtype=torch.bfloat16, device=d)
        self.dpre = torch.zeros(rows, self.H, dtype=torch.bfloat16, device=d)
        self.rng_base = None
        if init_params is not None:
            self.load_params(init_params)

    def load_params(self, params: Dict) -> None:
        """Loads a *global* reference-layout parameter tree (same values on every rank)."""
        self.load_flat_params({f"{a}/{b}": v for a, sub in params.items() for b, v in sub.items()})

    def params_tree(self, gather: bool = True) -> Dict:
        return self._nest(self.flat_params(gather=gather), _ORDER)

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
        xb = self.x_bf[:, : self.In]
        if self.x_bf.shape[1] == self.In:
            self._begin_step(self.x_in, self.x_bf)     # one launch
        else:
            self._begin_step(None, None)
            xb.copy_(self.x_in)
        self._gather_small()
        if self.fuse_minibatches:
            self._fwd_bwd(xb, self.labels, salts, mb, accumulate=False, last=True)
        else:
            for i in range(nmb):
                self._fwd_bwd(xb[i * mb:(i + 1) * mb], self.labels[i * mb:(i + 1) * mb], [salts[i]], mb,
                              accumulate=i > 0, last=i == nmb - 1)
        self._fold_end(next_rng)
        if self.overlap:
            torch.cuda.current_stream().wait_stream(self.side_stream)
            self._sync_bucket(1, last=True)
        else:
            if self.fork_wgrad:
                torch.cuda.current_stream().wait_stream(self.side_stream)     # join: output layer's weight gradient
            self._sync_all_buckets()
        self._end_step(next_rng)

    def _fwd_bwd(self, xs, lab, salts, rows_per_salt: int, *, accumulate: bool, last: bool) -> None:
        """Forward + backward of one row block (a minibatch, or all minibatches at once)."""
        logits = self.logits[:, : self.Cn]
        dlogits = self.dlogits[:, : self.Cn]
        det = self.deterministic
        sk = dict(split_k=1) if det else {}     # deterministic: no arrival-order sums (split-K, atomic column sums)
        db2 = None if det else self.grad1d("output_dense/bias")
        db1 = None if det else self.grad1d("input_dense/bias")
        drop = dict(drop_key=self.rng_base, drop_rate=self.rate, drop_salts=salts, drop_rows_per_salt=rows_per_salt)
        self.gemm_w(xs, "input_dense/kernel", self.h, bias=self.vec("input_dense/bias"), act="silu", pre=self.pre, **drop)
        w2_shard = self._weight("output_dense/kernel")[2]
        if self.fuse_xent and self.Cn <= 32 and not (w2_shard is not None and w2_shard[0]):     # not for a class-sharded head
            # head + softmax cross-entropy + its bias gradient in one launch: the logits stay in TMEM
            self.gemm_w(self.h, "output_dense/kernel", dlogits, bias=self.vec("output_dense/bias"), xent_labels=lab,
                        xent_metrics=self.step_metrics, xent_scale=1.0 / self.mb, dbias=db2)
        else:
            self.gemm_w(self.h, "output_dense/kernel", logits, bias=self.vec("output_dense/bias"))
            K.softmax_xent(logits, lab, grad_scale=1.0 / self.mb, dlogits=dlogits, metrics=self.step_metrics, dbias=db2)
        if det:
            self.grad1d("output_dense/bias").add_(dlogits.float().sum(0))
        if self.fork_wgrad:
            # dW2 only depends on h and dlogits: in a latency-bound step it runs as a parallel branch of
            # the graph next to the dgrad -> dW1 chain instead of in front of it
            main = torch.cuda.current_stream()
            self.side_stream.wait_stream(main)
            with torch.cuda.stream(self.side_stream):
                K.gemm(self.h, dlogits, self.grad2d("output_dense/kernel"), a_mn=True, b_mn=True, accumulate=accumulate, **sk)
        else:
            K.gemm(self.h, dlogits, self.grad2d("output_dense/kernel"), a_mn=True, b_mn=True, accumulate=accumulate, **sk)
        self.gemm_wT(dlogits, "output_dense/kernel", self.dpre, bwd=True, aux=self.pre, act="silu", dbias=db1, **drop)
        if det:
            self.grad1d("input_dense/bias").add_(self.dpre.float().sum(0))
        if last and self.overlap:
            # output layer's matrix: reduce + update under the rest of the backward. Queued only after
            # the dgrad above, the last reader of W2 in this step: the update rewrites W2 on every rank.
            self._sync_bucket_async(0)
        K.gemm(xs, self.dpre, self.grad2d("input_dense/kernel"), a_mn=True, b_mn=True, accumulate=accumulate, **sk)

"""Fused engine for GPT-2 (small / medium shapes): the transformer configurations of
BASELINE.json (data-parallel, FSDP with the all-gather fused into the GEMMs, 2-D data x fsdp mesh).

Forward and backward are an explicit schedule of our kernels over flat buffers:

    embedding -> [ LN -> GEMM(qkv,+bias) -> attention -> GEMM(proj,+bias,+residual)
                   -> LN -> GEMM(fc,+bias,GELU) -> GEMM(proj,+bias,+residual) ] x L
              -> LN -> GEMM(tied head) -> softmax-xent (in place: logits become dlogits)

Backward mirrors it; weight gradients accumulate straight into the flat fp32 gradient buffer
(``accumulate`` over minibatches = the reference's accum_grads, util.py:41-78), GELU' rides in the
dgrad epilogue, LayerNorm backward adds the residual-branch gradient. Each layer is one gradient
bucket: as soon as a layer's backward is queued its fused reduce+AdamW launch goes to the side
stream and overlaps the backward of the layers below it (DP: NVLS all-reduce, data_paral.py:210-217;
FSDP: reduce-scatter + sharded update, param_sharding.py:134-138,353).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch

from .. import rng as _rng
from ..mesh import Mesh
from ..models.gpt2 import padded_vocab
from ..ops import kernels as K
from ..state import Batch
from .base import FlatLayout, FusedEngine


def _layer_leaves(i: int, c: int, bucket: int):
    p = f"h_{i}/"
    return [
        (p + "ln_1/scale", (c,), bucket), (p + "ln_1/bias", (c,), bucket),
        (p + "attn/c_attn/kernel", (c, 3 * c), bucket), (p + "attn/c_attn/bias", (3 * c,), bucket),
        (p + "attn/c_proj/kernel", (c, c), bucket), (p + "attn/c_proj/bias", (c,), bucket),
        (p + "ln_2/scale", (c,), bucket), (p + "ln_2/bias", (c,), bucket),
        (p + "mlp/c_fc/kernel", (c, 4 * c), bucket), (p + "mlp/c_fc/bias", (4 * c,), bucket),
        (p + "mlp/c_proj/kernel", (4 * c, c), bucket), (p + "mlp/c_proj/bias", (c,), bucket),
    ]


class GPT2Engine(FusedEngine):
    def __init__(self, config, mesh: Mesh, *, mode: str = "dp", batch_per_rank: int, seq_len: Optional[int] = None,
                 lr: float = 3e-4, num_minibatches: int = 1, init_params: Optional[Dict] = None,
                 layer_ids: Optional[Sequence[int]] = None, has_embed: bool = True, has_head: bool = True,
                 slots: int = 1, **kw):
        """``layer_ids`` / ``has_embed`` / ``has_head`` / ``slots`` make the engine one *pipeline stage*
        (``engine.pp.PipelineEngine``): it owns only those transformer blocks (plus the embeddings on the
        first stage, the final norm and a tied copy of ``wte`` on the last), keeps ``slots`` sets of
        activations (one per microbatch in flight) and its forward / backward take and return the
        activations at the stage boundary instead of tokens and logits."""
        kw.setdefault("comm_blocks", 128)   # hundreds of MB per bucket: keep many bytes in flight
        tokens = (batch_per_rank // num_minibatches) * (seq_len or config.n_ctx)
        if kw.get("fsdp_weights", "auto") == "auto":   # many row tiles per GEMM: gather once per layer
            kw["fsdp_weights"] = "gather" if tokens > 1024 else "fused"
        super().__init__(mesh, mode=mode, lr=lr, num_minibatches=num_minibatches,
                         data_axis_name=config.get("data_axis_name", "data"), **kw)
        if self.deterministic:
            raise NotImplementedError(
                "GPT2Engine has no deterministic schedule: embedding_bwd scatters with atomics, attention backward "
                "reduce-adds dQ tiles in arrival order, LayerNorm backward and the GELU' dgrad add their column sums "
                "atomically, and the weight-gradient GEMMs use split-K; results are reproducible to fp32 rounding, "
                "not bit for bit")
        self.cfg = config
        self.pdl = tokens < 4096       # long persistent kernels: dependent launch only ties up SMs (see FusedEngine.pdl)
        self.layer_ids = list(range(config.n_layer)) if layer_ids is None else list(layer_ids)
        self.has_embed, self.has_head, self.slots = has_embed, has_head, slots
        self.is_stage = not (has_embed and has_head and len(self.layer_ids) == config.n_layer)
        self.L, self.Hh, self.C = len(self.layer_ids), config.n_head, config.d_model
        self.T = seq_len or config.n_ctx
        assert self.T % 128 == 0 and self.T <= config.n_ctx
        self.V = config.vocab_size
        self.Vp = padded_vocab(self.V)
        self.B = batch_per_rank
        assert self.B % num_minibatches == 0
        self.mbB = self.B // num_minibatches
        self.M = self.mbB * self.T
        c, L = self.C, self.L
        # Buckets in gradient-ready order: the matrices of layers L-1 .. 0 (bucket j = layer L-1-j), then
        # one last bucket with the embeddings and *all* 1-D leaves (biases, norm scales). Keeping the
        # vectors together makes their gradients one contiguous region: it is cleared with a single
        # launch per step and filled by atomics from the kernels that already hold the data (dgrad
        # epilogues, LayerNorm backward) instead of separate column-sum passes.
        mats, vecs = [], []
        for j, i in enumerate(reversed(self.layer_ids)):
            for path, shape, _ in _layer_leaves(i, c, j):
                (mats if len(shape) == 2 else vecs).append((path, shape, j if len(shape) == 2 else L))
        tables = ([("wte", (self.Vp, c), L)] if (has_embed or has_head) else []) + \
            ([("wpe", (config.n_ctx, c), L)] if has_embed else [])
        if has_head:
            vecs = vecs + [("ln_f/scale", (c,), L), ("ln_f/bias", (c,), L)]
        leaves = mats + tables + vecs
        self._setup_buffers(FlatLayout(leaves, self.world))
        if vecs:
            first_vec, last_vec = self.layout.leaves[vecs[0][0]], self.layout.leaves[vecs[-1][0]]
            self._vec_grads = self.g32[first_vec.offset: last_vec.offset + last_vec.ld]
        else:
            self._vec_grads = None
        d, M = self.device, self.M
        bf, f32 = torch.bfloat16, torch.float32
        self.tokens = torch.zeros(self.B, self.T, dtype=torch.int64, device=d)
        self.targets = torch.zeros(self.B, self.T, dtype=torch.int64, device=d)
        z = lambda *s, dt=bf: torch.zeros(*s, dtype=dt, device=d)
        def new_acts():
            return [dict(a=z(M, c), qkv=z(M, 3 * c), att=z(M, c), x1=z(M, c), m2=z(M, c), hpre=z(M, 4 * c),
                         hact=z(M, 4 * c), x2=z(M, c), mean1=z(M, dt=f32), rstd1=z(M, dt=f32),
                         mean2=z(M, dt=f32), rstd2=z(M, dt=f32), lse=z(self.mbB * self.Hh * self.T, dt=f32))
                    for _ in range(L)]
        self.slot_acts: List[List[Dict[str, torch.Tensor]]] = [new_acts() for _ in range(slots)]
        self.slot_x0 = [z(M, c) for _ in range(slots)]        # stage input: embeddings, or the activations received
        self.acts, self.x0 = self.slot_acts[0], self.slot_x0[0]
        if has_head:
            self.xf, self.meanf, self.rstdf = z(M, c), z(M, dt=f32), z(M, dt=f32)
            self.logits = z(M, self.Vp)
        if self.is_stage:
            self.slot_dx_top = [z(M, c) for _ in range(slots)]   # gradient entering the stage from above
            self.slot_dx_out = [z(M, c) for _ in range(slots)]   # gradient leaving the stage downwards
        # backward temporaries
        self.slot_dx_top0 = self.slot_dx_top[0] if self.is_stage else z(M, c)
        self.dxa, self.dxb = z(M, c), z(M, c)
        self.dh, self.dm = z(M, 4 * c), z(M, c)
        self.dqkv = z(M, 3 * c)
        self.att_scratch = z(K.attention_bwd_scratch_numel(self.mbB, self.T, self.Hh), dt=f32) if L else None
        self.rng_base = None
        if init_params is not None:
            self.load_params(init_params)

    def _override_shard_dim(self, path: str, shape, dim):
        # Embedding tables are consumed by row lookups: shard them on rows (or not at all). The
        # reference's rule picks rows anyway whenever the table is taller than wide, which holds for
        # every real configuration (50304 x C, 1024 x C).
        if path in ("wte", "wpe"):
            return 0 if shape[0] % self.S == 0 else None
        return dim

    # ------------------------------------------------------------------ parameters ------------
    @staticmethod
    def _flatten(tree: Dict, prefix: str = "") -> Dict[str, torch.Tensor]:
        out = {}
        for k, v in tree.items():
            if isinstance(v, dict):
                out.update(GPT2Engine._flatten(v, f"{prefix}{k}/"))
            else:
                out[f"{prefix}{k}"] = v
        return out

    def load_params(self, params: Dict) -> None:
        self.load_flat_params(self._flatten(params))

    def params_tree(self, gather: bool = True) -> Dict:
        flat = self.flat_params(gather=gather)
        order = ["wte", "wpe"] + [p for i in self.layer_ids for p, _, _ in _layer_leaves(i, self.C, 0)] + ["ln_f/scale", "ln_f/bias"]
        return self._nest(flat, [p for p in order if p in flat])

    # ------------------------------------------------------------------ step ------------------
    def _load_batch(self, batch: Batch) -> None:
        self.tokens.copy_(batch.inputs, non_blocking=True)
        self.targets.copy_(batch.labels, non_blocking=True)

    def _rows_per_step(self) -> int:
        return self.B * self.T * self.world

    def _embed(self, tok: torch.Tensor, out: Optional[torch.Tensor] = None) -> None:
        out = self.x0 if out is None else out
        if self.mode == "fsdp" and self.S > 1 and (self.shard_dim["wte"] is not None or self.shard_dim["wpe"] is not None):
            K.embedding_fwd_sharded(tok, self._table_peers("wte"), self._table_peers("wpe"), out, self.C)
        else:
            K.embedding_fwd(tok, self._table("wte"), self._table("wpe"), out)

    def _table(self, path: str) -> torch.Tensor:
        return self.layout.view2d(self.pb, path) if self.mode == "dp" else self._local_view(self.pb, path)

    def _table_peers(self, path: str):
        r, c, ld = self._local_shape2d(path)
        if self.shard_dim[path] is None:
            return [self._local_view(self.pb, path).data_ptr()], 1 << 30
        return ([self._peer_base(g) + self.arena.offsets["bf16"] + self.loc_off[path] * 2
                 for g in self.shard_peer_ranks], r)

    def _step_body(self) -> None:
        nmb = self.num_minibatches
        self._begin_step(None, None)   # metrics + every bias / norm gradient cleared, peers' last update awaited: one launch
        self._gather_small()
        next_rng = _rng.split(_rng.Key(self.rng_base, 0), 2)[0]
        for i in range(nmb):
            sl = slice(i * self.mbB, (i + 1) * self.mbB)
            self._forward(self.tokens[sl])
            self._backward(self.tokens[sl], self.targets[sl], accumulate=i > 0, last=i == nmb - 1)
        self._fold_end(next_rng)
        if self.overlap:
            torch.cuda.current_stream().wait_stream(self.side_stream)
            self._sync_bucket(self.L, last=True)
        else:
            self._sync_all_buckets()
        self._end_step(next_rng)

    def _forward(self, tok: Optional[torch.Tensor], slot: int = 0) -> torch.Tensor:
        """Forward of one microbatch through this engine's blocks. The input is ``tok`` (engines that own
        the embeddings) or whatever the caller put into ``slot_x0[slot]`` (a pipeline stage's received
        activations). Returns the output activations (the logits stay in ``self.logits`` when the
        engine owns the head)."""
        acts, x = self.slot_acts[slot], self.slot_x0[slot]
        if self.has_embed:
            self._embed(tok, x)
        L = self.L
        bucket_of = lambda li: L - 1 - li
        if L:
            self.prefetch_bucket(bucket_of(0))
        for li, i in enumerate(self.layer_ids):
            p, A = f"h_{i}/", acts[li]
            self.prefetch_bucket(bucket_of(li + 1) if li + 1 < L else L, keep=bucket_of(li))   # next layer / the tied head
            K.layernorm_fwd(x, self.vec(p + "ln_1/scale"), self.vec(p + "ln_1/bias"), A["a"], A["mean1"], A["rstd1"])
            self.gemm_w(A["a"], p + "attn/c_attn/kernel", A["qkv"], bias=self.vec(p + "attn/c_attn/bias"))
            K.attention_fwd(A["qkv"], A["att"], A["lse"], self.mbB, self.T, self.Hh)
            self.gemm_w(A["att"], p + "attn/c_proj/kernel", A["x1"], bias=self.vec(p + "attn/c_proj/bias"), resid=x)
            K.layernorm_fwd(A["x1"], self.vec(p + "ln_2/scale"), self.vec(p + "ln_2/bias"), A["m2"], A["mean2"], A["rstd2"])
            self.gemm_w(A["m2"], p + "mlp/c_fc/kernel", A["hact"], bias=self.vec(p + "mlp/c_fc/bias"), act="gelu",
                        pre=A["hpre"])
            self.gemm_w(A["hact"], p + "mlp/c_proj/kernel", A["x2"], bias=self.vec(p + "mlp/c_proj/bias"), resid=A["x1"])
            A["x_in"] = x
            x = A["x2"]
        if self.has_head:
            K.layernorm_fwd(x, self.vec("ln_f/scale"), self.vec("ln_f/bias"), self.xf, self.meanf, self.rstdf)
            self.gemm_wT(self.xf, "wte", self.logits)          # tied head: logits = xf @ wte^T
        self._x_last = x
        return x

    def _head_backward(self, tgt: torch.Tensor, dx_out: torch.Tensor, *, accumulate: bool) -> None:
        """Loss + backward of the final norm and the tied head for the microbatch whose logits are in
        ``self.logits``: writes the gradient of the residual stream into ``dx_out``."""
        g2, g1 = self.grad2d, self.grad1d
        logits = self.logits[:, : self.V]
        K.softmax_xent_bf16_(logits, tgt.reshape(-1), grad_scale=1.0 / self.M, metrics=self.step_metrics)
        dlog = self.logits                                   # now the gradient (pad columns stay zero)
        if self.L:
            self.prefetch_bucket(0, keep=self.L)             # top layer's weights while the head runs
        K.gemm(dlog, self.xf, g2("wte"), a_mn=True, b_mn=True, accumulate=accumulate)      # dwte = dlogits^T xf
        self.gemm_w(dlog, "wte", self.dm)                                                   # dxf = dlogits wte
        # d(residual stream) out of the final norm; its column sums are the top layer's mlp/c_proj bias gradient
        top_bias = g1(f"h_{self.layer_ids[-1]}/mlp/c_proj/bias") if self.L else None
        K.layernorm_bwd(self.dm, self._x_last, self.vec("ln_f/scale"), self.meanf, self.rstdf, dx_out, g1("ln_f/scale"),
                        g1("ln_f/bias"), dxsum=top_bias)

    def _layers_backward(self, dx: torch.Tensor, *, accumulate: bool, last: bool, slot: int = 0,
                         final_out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Backward through this engine's blocks, top to bottom, for the microbatch saved in ``slot``.
        ``dx`` is the gradient of the output activations (read only). Returns the gradient of the input
        activations (in ``final_out`` if given)."""
        g2, g1 = self.grad2d, self.grad1d
        acc, L = accumulate, self.L
        acts = self.slot_acts[slot]
        for j, li in enumerate(range(L - 1, -1, -1)):
            i = self.layer_ids[li]
            p, A = f"h_{i}/", acts[li]
            if li > 0:
                self.prefetch_bucket(j + 1, keep=j)          # the layer below, needed next
            # x2 = x1 + gelu(m2 Wfc + bfc) Wproj + bproj          (bias gradient of Wproj: already summed from dx)
            K.gemm(A["hact"], dx, g2(p + "mlp/c_proj/kernel"), a_mn=True, b_mn=True, accumulate=acc)
            self.gemm_wT(dx, p + "mlp/c_proj/kernel", self.dh, bwd=True, aux=A["hpre"], act="gelu",
                         dbias=g1(p + "mlp/c_fc/bias"))      # dh = d(pre-GELU); its column sums = c_fc bias gradient
            K.gemm(A["m2"], self.dh, g2(p + "mlp/c_fc/kernel"), a_mn=True, b_mn=True, accumulate=acc)
            self.gemm_wT(self.dh, p + "mlp/c_fc/kernel", self.dm)
            K.layernorm_bwd(self.dm, A["x1"], self.vec(p + "ln_2/scale"), A["mean2"], A["rstd2"], self.dxa,
                            g1(p + "ln_2/scale"), g1(p + "ln_2/bias"), dresid=dx, dxsum=g1(p + "attn/c_proj/bias"))
            dx = self.dxa                                    # dx = d(x1)
            # x1 = x + att Wo + bo
            K.gemm(A["att"], dx, g2(p + "attn/c_proj/kernel"), a_mn=True, b_mn=True, accumulate=acc)
            self.gemm_wT(dx, p + "attn/c_proj/kernel", self.dm)                      # d(att)
            K.attention_bwd(A["qkv"], A["att"], self.dm, A["lse"], self.dqkv, self.att_scratch, self.mbB, self.T, self.Hh)
            K.gemm(A["a"], self.dqkv, g2(p + "attn/c_attn/kernel"), a_mn=True, b_mn=True, accumulate=acc)
            K.colsum(self.dqkv, g1(p + "attn/c_attn/bias"), accumulate=True)
            self.gemm_wT(self.dqkv, p + "attn/c_attn/kernel", self.dm)               # d(a)
            below = g1(f"h_{self.layer_ids[li - 1]}/mlp/c_proj/bias") if li > 0 else None
            out = final_out if (li == 0 and final_out is not None) else self.dxb
            K.layernorm_bwd(self.dm, A["x_in"], self.vec(p + "ln_1/scale"), A["mean1"], A["rstd1"], out,
                            g1(p + "ln_1/scale"), g1(p + "ln_1/bias"), dresid=dx, dxsum=below)
            dx = out                                         # dx = d(x_in)
            if last and self.overlap:
                self._sync_bucket_async(j)                   # this layer's matrices: reduce + update under the rest
        return dx

    def _embed_backward(self, tok: torch.Tensor, dx: torch.Tensor, *, accumulate: bool, zero_wte: bool = False) -> None:
        if not accumulate:
            K.scale_(self.grad2d_full("wpe"), 0.0)
            if zero_wte:          # a first pipeline stage: no head GEMM has written the table's gradient
                K.scale_(self.grad2d_full("wte"), 0.0)
        K.embedding_bwd(tok, dx, self.grad2d_full("wte"), self.grad2d_full("wpe"))

    def _backward(self, tok: torch.Tensor, tgt: torch.Tensor, *, accumulate: bool, last: bool) -> None:
        self._head_backward(tgt, self.slot_dx_top0, accumulate=accumulate)
        dx = self._layers_backward(self.slot_dx_top0, accumulate=accumulate, last=last)
        self._embed_backward(tok, dx, accumulate=accumulate)

    def grad2d_full(self, path: str) -> torch.Tensor:
        """Contiguous [rows, ld] gradient region of a 2-D leaf whose last dim needs no padding."""
        sp = self.layout.leaves[path]
        assert sp.ld == sp.cols
        return self.g32[sp.offset: sp.offset + sp.rows * sp.ld].view(sp.rows, sp.cols)

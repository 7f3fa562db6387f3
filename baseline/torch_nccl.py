"""Stock-PyTorch data-parallel / FSDP-style / pipeline train steps for the MLP and GPT-2 workloads:
library GEMMs (cuBLAS), SDPA attention, NCCL collectives, torch's fused AdamW. This is the comparison
arm of BASELINE.md section 4 ("a path that only calls NCCL is the baseline, not the product").

DP   : replicated fp32 params, bf16 autocast compute, gradient accumulation over minibatches,
       ``DistributedDataParallel`` (25 MB buckets reduced while the backward still runs, no_sync for all
       but the last minibatch) for models with more than one bucket, one flat ``all_reduce`` for the
       four-leaf MLP; ``torch.optim.AdamW(fused=True)`` (the dataflow of data_paral.py:193-238).
       With ``graph=True`` the whole step (forward, backward, all-reduce, optimizer) is captured in a
       CUDA graph when the path allows it (no DDP hooks): what XLA's single compiled step would be.
FSDP : each parameter stored as a 1/N flat shard; before the forward every layer's weights are
       materialised with ``all_gather_into_tensor``; after the backward the full gradients go
       through ``reduce_scatter_tensor`` (mean) into the shard, AdamW updates the shards
       (the dataflow of param_sharding.py:129-175,343-367).
PP   : GPipe over ``stages`` ranks: each stage runs its blocks with autograd per microbatch, activations
       and their gradients travel with ``batch_isend_irecv``; same layer split as the product path.
"""
from __future__ import annotations

from typing import List

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


class MLP(nn.Module):
    def __init__(self, inp=784, hidden=512, classes=10, p=0.1):
        super().__init__()
        self.input_dense = nn.Linear(inp, hidden)
        self.output_dense = nn.Linear(hidden, classes)
        self.p = p

    def forward(self, x):
        h = F.dropout(F.silu(self.input_dense(x)), self.p, self.training)
        return self.output_dense(h).float()


class Block(nn.Module):
    def __init__(self, c, h):
        super().__init__()
        self.ln_1, self.ln_2 = nn.LayerNorm(c), nn.LayerNorm(c)
        self.c_attn, self.c_proj = nn.Linear(c, 3 * c), nn.Linear(c, c)
        self.c_fc, self.mlp_proj = nn.Linear(c, 4 * c), nn.Linear(4 * c, c)
        self.h = h

    def forward(self, x):
        b, t, c = x.shape
        q, k, v = self.c_attn(self.ln_1(x)).split(c, dim=2)
        q, k, v = (z.view(b, t, self.h, c // self.h).transpose(1, 2) for z in (q, k, v))
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(b, t, c)
        x = x + self.c_proj(y)
        return x + self.mlp_proj(F.gelu(self.c_fc(self.ln_2(x)), approximate="tanh"))


class GPT2(nn.Module):
    def __init__(self, n_layer=12, n_head=12, c=768, vocab=50304, n_ctx=1024):
        super().__init__()
        self.wte, self.wpe = nn.Embedding(vocab, c), nn.Embedding(n_ctx, c)
        self.h = nn.ModuleList([Block(c, n_head) for _ in range(n_layer)])
        self.ln_f = nn.LayerNorm(c)

    def forward(self, tok):
        x = self.wte(tok) + self.wpe(torch.arange(tok.shape[1], device=tok.device))
        for blk in self.h:
            x = blk(x)
        return F.linear(self.ln_f(x), self.wte.weight).float()


class BaselineTrainer:
    """``step(inputs, labels) -> loss`` with DP, FSDP-style or GPipe communication over NCCL."""

    def __init__(self, model: nn.Module, *, mode: str, lr: float, minibatches: int, world: int, rank: int,
                 stages: int = 1, graph: bool = False):
        self.model, self.mode, self.nmb, self.world, self.rank = model, mode, minibatches, world, rank
        self.stages = stages
        self.graph, self._g, self._calls = None, graph, 0
        if mode == "pp":
            self._setup_pp(lr)
            return
        self.params: List[torch.Tensor] = [p for p in model.parameters()]
        self.ddp = None
        capt = dict(capturable=True) if graph else {}
        if mode == "fsdp" and world > 1:
            # flat shards (padded) are the source of truth; module params are (re)filled by all-gather
            self.full_numel = [p.numel() for p in self.params]
            self.pad = [(-n) % world for n in self.full_numel]
            self.shards = []
            for p, pad in zip(self.params, self.pad):
                flat = F.pad(p.data.reshape(-1), (0, pad))
                self.shards.append(flat.chunk(world)[rank].clone().requires_grad_(False))
            self.opt = torch.optim.AdamW(self.shards, lr=lr, weight_decay=1e-4, eps=1e-8, fused=True, **capt)
            for s in self.shards:
                s.grad = torch.zeros_like(s)
        else:
            self.opt = torch.optim.AdamW(self.params, lr=lr, weight_decay=1e-4, eps=1e-8, fused=True, **capt)
            nbytes = sum(p.numel() * 4 for p in self.params)
            if mode == "dp" and world > 1 and nbytes > (25 << 20):
                from torch.nn.parallel import DistributedDataParallel as DDP
                self.ddp = DDP(model, bucket_cap_mb=25, gradient_as_bucket_view=True)
                self._g = False          # DDP's reducer hooks are not captured: eager step, overlapped buckets
        self.flat_grad = None

    # ------------------------------------------------------------------ DP / FSDP ---------------------
    def _gather_params(self):
        for p, s, n, pad in zip(self.params, self.shards, self.full_numel, self.pad):
            full = torch.empty(s.numel() * self.world, dtype=s.dtype, device=s.device)
            dist.all_gather_into_tensor(full, s)
            p.data.copy_(full[:n].view_as(p))

    def _step_impl(self, x, y):
        import contextlib
        model, nmb = (self.ddp or self.model), self.nmb
        if self.mode == "fsdp" and self.world > 1:
            self._gather_params()
        for p in self.params:
            p.grad = None
        total = 0.0
        mb = x.shape[0] // nmb
        for i in range(nmb):
            sync = contextlib.nullcontext() if (self.ddp is None or i == nmb - 1) else self.ddp.no_sync()
            with sync:
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    logits = model(x[i * mb:(i + 1) * mb])
                loss = F.cross_entropy(logits.reshape(-1, logits.shape[-1]), y[i * mb:(i + 1) * mb].reshape(-1)) / nmb
                loss.backward()
            total = total + loss.detach()
        if self.world > 1 and self.ddp is None:
            if self.mode == "dp":
                flat = torch.cat([p.grad.reshape(-1) for p in self.params])
                dist.all_reduce(flat)
                flat /= self.world
                off = 0
                for p in self.params:
                    p.grad.copy_(flat[off: off + p.numel()].view_as(p))
                    off += p.numel()
            else:
                for p, s, pad in zip(self.params, self.shards, self.pad):
                    g = F.pad(p.grad.reshape(-1), (0, pad))
                    dist.reduce_scatter_tensor(s.grad, g)
                    s.grad /= self.world
        self.opt.step()
        return total

    def step(self, x, y):
        if self.mode == "pp":
            return self._step_pp(x, y)
        if not self._g:
            return self._step_impl(x, y)
        # CUDA-graphed step: three eager steps (lazy initialisation), then one capture, then replays
        if self.graph is None:
            self._calls += 1
            if self._calls <= 3:
                return self._step_impl(x, y)
            try:
                self._sx, self._sy = x.clone(), y.clone()
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, capture_error_mode="thread_local"):
                    self._sloss = self._step_impl(self._sx, self._sy)
                self.graph = g
            except Exception as e:  # noqa: BLE001  (some paths cannot be captured: stay eager, say so once)
                print(f"[baseline] CUDA graph capture failed ({type(e).__name__}: {str(e)[:120]}); running eagerly", flush=True)
                self._g = False
                torch.cuda.synchronize()
                return self._step_impl(x, y)
        self._sx.copy_(x, non_blocking=True)
        self._sy.copy_(y, non_blocking=True)
        self.graph.replay()
        return self._sloss

    # ------------------------------------------------------------------ GPipe ---------------------------
    def _setup_pp(self, lr: float) -> None:
        """The model's blocks are dealt to ``stages`` consecutive ranks (same cost-balanced split as the
        product path); rank r is stage r % stages of data replica r // stages."""
        from jax_distributed_tuts_b200.engine.pp import balanced_split
        m, S = self.model, self.stages
        self.s = self.rank % S
        self.first, self.last = self.s == 0, self.s == S - 1
        c = m.wte.weight.shape[1]
        split = balanced_split(len(m.h), S, m.wte.weight.shape[0] / (12.0 * c))
        self.blocks = nn.ModuleList([m.h[i] for i in split[self.s]])
        ps = [p for b in self.blocks for p in b.parameters()]
        if self.first:
            ps += [m.wte.weight, m.wpe.weight]
        if self.last:
            ps += list(m.ln_f.parameters()) + ([m.wte.weight] if not self.first else [])
        self.params = ps
        self.opt = torch.optim.AdamW(ps, lr=lr, weight_decay=1e-4, eps=1e-8, fused=True)
        self.dp_group = None
        if self.world > S:
            for st in range(S):
                g = dist.new_group([st + S * d for d in range(self.world // S)])
                if st == self.s:
                    self.dp_group = g

    def _stage_forward(self, x):
        m = self.model
        if self.first:
            x = m.wte(x) + m.wpe(torch.arange(x.shape[1], device=x.device))
        for blk in self.blocks:
            x = blk(x)
        if self.last:
            x = F.linear(m.ln_f(x), m.wte.weight).float()
        return x

    @staticmethod
    def _p2p(send, dst, recv, src):
        ops = []
        if send is not None:
            ops.append(dist.P2POp(dist.isend, send, dst))
        if recv is not None:
            ops.append(dist.P2POp(dist.irecv, recv, src))
        for w in dist.batch_isend_irecv(ops):
            w.wait()

    def _step_pp(self, x, y):
        S, nmb, m = self.stages, self.nmb, self.model
        base = self.rank - self.s
        prv, nxt = base + self.s - 1, base + self.s + 1
        for p in self.params:
            p.grad = None
        mb = x.shape[0] // nmb
        c = m.wte.weight.shape[1]
        saved, total = [], torch.zeros((), device=x.device)
        for i in range(nmb):
            if self.first:
                inp, xin = x[i * mb:(i + 1) * mb], None
            else:
                xin = torch.empty(mb, x.shape[1], c, dtype=torch.bfloat16, device=x.device)
                self._p2p(None, None, xin, prv)
                xin.requires_grad_(True)
                inp = xin
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = self._stage_forward(inp)
            if self.last:
                loss = F.cross_entropy(out.reshape(-1, out.shape[-1]), y[i * mb:(i + 1) * mb].reshape(-1)) / nmb
                total = total + loss.detach()
                saved.append((xin, loss))
            else:
                self._p2p(out.detach().to(torch.bfloat16).contiguous(), nxt, None, None)
                saved.append((xin, out))
        for i in range(nmb - 1, -1, -1):
            xin, out = saved[i]
            if self.last:
                out.backward()
            else:
                g = torch.empty_like(out, dtype=torch.bfloat16)
                self._p2p(None, None, g, nxt)
                out.backward(g.to(out.dtype))
            if xin is not None:
                self._p2p(xin.grad.to(torch.bfloat16).contiguous(), prv, None, None)
        if S > 1 and (self.first or self.last):       # tied embedding: both ends keep the summed gradient
            g = m.wte.weight.grad
            other = torch.empty_like(g)
            peer = base + (S - 1 if self.first else 0)
            self._p2p(g.contiguous(), peer, other, peer)
            g.add_(other)
        if self.dp_group is not None:
            for p in self.params:
                dist.all_reduce(p.grad, group=self.dp_group)
                p.grad /= (self.world // S)
        self.opt.step()
        if self.world > 1:
            dist.all_reduce(total)       # only last stages contribute
            total = total / max(1, self.world // S)
        return total

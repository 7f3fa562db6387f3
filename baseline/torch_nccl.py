"""Stock-PyTorch data-parallel / FSDP-style train steps for the MLP and GPT-2 workloads.

DP   : replicated fp32 params, bf16 autocast compute, gradient accumulation over minibatches,
       one flat ``all_reduce`` of the gradients (mean), ``torch.optim.AdamW(fused=True)``
       (the dataflow of data_paral.py:193-238).
FSDP : each parameter stored as a 1/N flat shard; before the forward every layer's weights are
       materialised with ``all_gather_into_tensor``; after the backward the full gradients go
       through ``reduce_scatter_tensor`` (mean) into the shard, AdamW updates the shards
       (the dataflow of param_sharding.py:129-175,343-367).
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
    """``step(inputs, labels) -> loss`` with DP or FSDP-style communication over NCCL."""

    def __init__(self, model: nn.Module, *, mode: str, lr: float, minibatches: int, world: int, rank: int):
        self.model, self.mode, self.nmb, self.world, self.rank = model, mode, minibatches, world, rank
        self.params: List[torch.Tensor] = [p for p in model.parameters()]
        if mode == "fsdp" and world > 1:
            # flat shards (padded) are the source of truth; module params are (re)filled by all-gather
            self.full_numel = [p.numel() for p in self.params]
            self.pad = [(-n) % world for n in self.full_numel]
            self.shards = []
            for p, pad in zip(self.params, self.pad):
                flat = F.pad(p.data.reshape(-1), (0, pad))
                self.shards.append(flat.chunk(world)[rank].clone().requires_grad_(False))
            self.opt = torch.optim.AdamW(self.shards, lr=lr, weight_decay=1e-4, eps=1e-8, fused=True)
            for s in self.shards:
                s.grad = torch.zeros_like(s)
        else:
            self.opt = torch.optim.AdamW(self.params, lr=lr, weight_decay=1e-4, eps=1e-8, fused=True)
        self.flat_grad = None

    def _gather_params(self):
        for p, s, n, pad in zip(self.params, self.shards, self.full_numel, self.pad):
            full = torch.empty(s.numel() * self.world, dtype=s.dtype, device=s.device)
            dist.all_gather_into_tensor(full, s)
            p.data.copy_(full[:n].view_as(p))

    def step(self, x, y):
        model, nmb = self.model, self.nmb
        if self.mode == "fsdp" and self.world > 1:
            self._gather_params()
        for p in self.params:
            p.grad = None
        total = 0.0
        mb = x.shape[0] // nmb
        for i in range(nmb):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                logits = model(x[i * mb:(i + 1) * mb])
            loss = F.cross_entropy(logits.reshape(-1, logits.shape[-1]), y[i * mb:(i + 1) * mb].reshape(-1)) / nmb
            loss.backward()
            total = total + loss.detach()
        if self.world > 1:
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

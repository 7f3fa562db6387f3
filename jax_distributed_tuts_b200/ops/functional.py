"""Differentiable ops for the generic (pytree + autograd) tier.

On CUDA tensors every op runs our sm_100a kernels through ``torch.autograd.Function``; on CPU
tensors the same maths is written in plain PyTorch with bf16 rounding at the same points (this
is the oracle the GPU tests compare against, and what the gloo tier runs). There is no silent
fallback: a CUDA tensor with the library missing raises.

The fused engines (``engine/``) call the kernels directly and fuse further (activation
backward inside the dgrad epilogue, wgrad accumulation into the gradient bucket).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from .. import rng as _rng
from . import kernels as K


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


def _act_ref(x: torch.Tensor, act: Optional[str]) -> torch.Tensor:
    if act == "silu":
        return F.silu(x)
    if act == "gelu":
        return F.gelu(x, approximate="tanh")
    return x


def to_bf16_padded(x: torch.Tensor) -> torch.Tensor:
    """bf16 copy of a 1-D/2-D tensor whose row pitch is a multiple of 8 elements (TMA needs
    16-byte pitches); returns the logical view."""
    if x.dim() == 1:
        buf = torch.zeros(_pad8(x.shape[0]), dtype=torch.bfloat16, device=x.device)
        buf[: x.shape[0]].copy_(x)
        return buf[: x.shape[0]]
    rows, cols = x.shape
    if cols % 8 == 0 and x.is_contiguous():
        if x.dtype == torch.bfloat16:
            return x
        out = torch.empty(rows, cols, dtype=torch.bfloat16, device=x.device)
        if x.dtype == torch.float32 and x.is_cuda:
            return K.cast_f32_bf16(x, out)
        return out.copy_(x)
    buf = torch.zeros(rows, _pad8(cols), dtype=torch.bfloat16, device=x.device)
    view = buf[:, :cols]
    view.copy_(x)
    return view


class _DenseCUDA(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, kernel, bias, act, drop_rate, drop_key, drop_salt, out_fp32):
        lead = x.shape[:-1]
        x2 = to_bf16_padded(x.reshape(-1, x.shape[-1]))
        wb = to_bf16_padded(kernel)
        bb = to_bf16_padded(bias) if bias is not None else None
        m, n = x2.shape[0], kernel.shape[1]
        need_pre = act is not None or drop_rate > 0.0
        out_buf = torch.empty(m, _pad8(n), dtype=torch.float32 if out_fp32 else torch.bfloat16, device=x.device)
        out = out_buf[:, :n]
        pre = torch.empty(m, _pad8(n), dtype=torch.bfloat16, device=x.device)[:, :n] if need_pre else None
        K.gemm(x2, wb, out, b_mn=True, bias=bb, act=act, pre=pre, drop_key=drop_key, drop_salt=drop_salt,
               drop_rate=drop_rate)
        ctx.save_for_backward(x2, wb, pre, drop_key)
        ctx.meta = (act, drop_rate, drop_salt, x.dtype, kernel.dtype, bias is not None, lead, x.shape[-1])
        return out.reshape(*lead, n)

    @staticmethod
    def backward(ctx, dy):
        x2, wb, pre, drop_key = ctx.saved_tensors
        act, drop_rate, drop_salt, x_dtype, w_dtype, has_bias, lead, k = ctx.meta
        n = wb.shape[1]
        dy2 = to_bf16_padded(dy.reshape(-1, n))
        if pre is not None:
            dpre = torch.empty(dy2.shape[0], _pad8(n), dtype=torch.bfloat16, device=dy.device)[:, :n]
            K.act_dropout_bwd(pre, dy2, dpre, act, drop_key, drop_salt, drop_rate)
        else:
            dpre = dy2
        dw = torch.zeros(k, _pad8(n), dtype=torch.float32, device=dy.device)[:, :n]
        K.gemm(x2, dpre, dw, a_mn=True, b_mn=True)
        db = None
        if has_bias:
            db = torch.empty(n, dtype=torch.float32, device=dy.device)
            K.colsum(dpre, db)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(dy2.shape[0], _pad8(k), dtype=torch.bfloat16, device=dy.device)[:, :k]
            K.gemm(dpre, wb, dx)  # dy @ W^T : B stored [N'=k][K'=n]
            dx = dx.reshape(*lead, k).to(x_dtype)
        return dx, dw.to(w_dtype), db, None, None, None, None, None


def dense(x: torch.Tensor, kernel: torch.Tensor, bias: Optional[torch.Tensor] = None, *,
          act: Optional[str] = None, drop_rate: float = 0.0, drop_key: Optional[_rng.Key] = None,
          out_fp32: bool = False) -> torch.Tensor:
    """``dropout(act(x @ kernel + bias))`` with bf16 compute and fp32 accumulation.

    ``kernel`` is (in, out) like ``flax.linen.Dense`` (data_paral.py:86-98). Output is bf16
    unless ``out_fp32``. Dropout uses the counter-based mask of ``rng.dropout_mask`` over the
    flattened [rows, out] activation, so CPU and GPU agree bit for bit on which units drop.
    """
    if drop_rate > 0.0 and drop_key is None:
        raise ValueError("dropout needs an RNG key")
    if x.is_cuda:
        key_t = drop_key.base if drop_key is not None else None
        salt = _rng.salt_word(drop_key) if drop_key is not None else 0
        return _DenseCUDA.apply(x, kernel, bias, act, float(drop_rate), key_t, salt, out_fp32)
    # ---- CPU oracle: same rounding points as the kernel ----
    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1]).to(torch.bfloat16)
    y = x2.float() @ kernel.to(torch.bfloat16).float()
    if bias is not None:
        y = y + bias.to(torch.bfloat16).float()
    if act is not None or drop_rate > 0.0:
        y = y.to(torch.bfloat16).float()  # the saved pre-activation is bf16
    y = _act_ref(y, act)
    if drop_rate > 0.0:
        mask = _rng.dropout_mask(_rng.materialize(drop_key), y.shape[0], y.shape[1], drop_rate, device=y.device)
        y = torch.where(mask, y / (1.0 - drop_rate), torch.zeros_like(y))
    y = y if out_fp32 else y.to(torch.bfloat16)
    return y.reshape(*lead, kernel.shape[1])


class _XentCUDA(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        rows, classes = logits.shape
        lg = logits.float().contiguous()
        dl = torch.empty(rows, _pad8(classes), dtype=torch.bfloat16, device=logits.device)[:, :classes]
        row_loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
        metrics = torch.zeros(2, dtype=torch.float32, device=logits.device)
        K.softmax_xent(lg, labels, grad_scale=1.0, dlogits=dl, metrics=metrics, row_loss=row_loss)
        ctx.save_for_backward(dl)
        ctx.correct = metrics[1]
        ctx.in_dtype = logits.dtype
        return row_loss, metrics[1].detach()

    @staticmethod
    def backward(ctx, g_loss, _g_correct):
        (dl,) = ctx.saved_tensors
        return (dl.float() * g_loss[:, None]).to(ctx.in_dtype), None


def softmax_cross_entropy_with_integer_labels(logits: torch.Tensor, labels: torch.Tensor,
                                              return_correct: bool = False):
    """Per-row ``logsumexp(logits) - logits[label]`` (optax; data_paral.py:181). With
    ``return_correct`` also returns the number of rows whose argmax equals the label
    (data_paral.py:183), which the CUDA kernel produces in the same pass."""
    if logits.is_cuda:
        loss, correct = _XentCUDA.apply(logits, labels)
    else:
        lg = logits.float()
        loss = torch.logsumexp(lg, dim=-1) - lg.gather(-1, labels[:, None]).squeeze(-1)
        correct = (lg.argmax(-1) == labels).sum().to(torch.float32)
    return (loss, correct) if return_correct else loss


# ------------------------------------------------------------------------------------------------
# Transformer ops (GPT-2 configurations of BASELINE.json; not present in the reference)
# ------------------------------------------------------------------------------------------------
class _LayerNormCUDA(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, scale, bias, eps):
        lead, c = x.shape[:-1], x.shape[-1]
        x2 = x.reshape(-1, c).to(torch.bfloat16).contiguous()
        g, b = scale.to(torch.bfloat16).contiguous(), bias.to(torch.bfloat16).contiguous()
        y = torch.empty_like(x2)
        mean = torch.empty(x2.shape[0], dtype=torch.float32, device=x.device)
        rstd = torch.empty_like(mean)
        K.layernorm_fwd(x2, g, b, y, mean, rstd, eps)
        ctx.save_for_backward(x2, g, mean, rstd)
        ctx.meta = (lead, c, x.dtype, scale.dtype)
        return y.reshape(*lead, c)

    @staticmethod
    def backward(ctx, dy):
        x2, g, mean, rstd = ctx.saved_tensors
        lead, c, xdt, pdt = ctx.meta
        dy2 = dy.reshape(-1, c).to(torch.bfloat16).contiguous()
        dx = torch.empty_like(x2)
        dg = torch.zeros(c, dtype=torch.float32, device=dy.device)
        db = torch.zeros(c, dtype=torch.float32, device=dy.device)
        K.layernorm_bwd(dy2, x2, g, mean, rstd, dx, dg, db)
        return dx.reshape(*lead, c).to(xdt), dg.to(pdt), db.to(pdt), None


def layer_norm(x: torch.Tensor, scale: torch.Tensor, bias: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    """LayerNorm over the last dim; bf16 in/out, fp32 statistics."""
    if x.is_cuda:
        return _LayerNormCUDA.apply(x, scale, bias, eps)
    xf = x.to(torch.bfloat16).float()
    y = F.layer_norm(xf, (x.shape[-1],), scale.to(torch.bfloat16).float(), bias.to(torch.bfloat16).float(), eps)
    return y.to(torch.bfloat16)


class _AttentionCUDA(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, heads):
        b, t, c3 = qkv.shape
        q2 = qkv.to(torch.bfloat16).contiguous()
        out = torch.empty(b, t, c3 // 3, dtype=torch.bfloat16, device=qkv.device)
        lse = torch.empty(b, heads, t, dtype=torch.float32, device=qkv.device)
        K.attention_fwd(q2, out, lse, b, t, heads)
        ctx.save_for_backward(q2, out, lse)
        ctx.meta = (b, t, heads, qkv.dtype)
        return out

    @staticmethod
    def backward(ctx, dout):
        q2, out, lse = ctx.saved_tensors
        b, t, heads, dt = ctx.meta
        dqkv = torch.empty_like(q2)
        scratch = torch.empty(K.attention_bwd_scratch_numel(b, t, heads), dtype=torch.float32, device=dout.device)
        K.attention_bwd(q2, out, dout.to(torch.bfloat16).contiguous(), lse, dqkv, scratch, b, t, heads)
        return dqkv.to(dt), None


def causal_self_attention(qkv: torch.Tensor, heads: int) -> torch.Tensor:
    """qkv [batch, seq, 3*C] laid out [q | k | v], each [heads, 64] -> [batch, seq, C]. head_dim 64."""
    b, t, c3 = qkv.shape
    c = c3 // 3
    assert c == heads * 64, "the attention kernel is specialised for head_dim 64"
    if qkv.is_cuda:
        return _AttentionCUDA.apply(qkv, heads)
    q, k, v = (z.reshape(b, t, heads, 64).permute(0, 2, 1, 3).to(torch.bfloat16).float() for z in qkv.split(c, dim=-1))
    o = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    return o.permute(0, 2, 1, 3).reshape(b, t, c).to(torch.bfloat16)


class _EmbeddingCUDA(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tokens, wte, wpe):
        b, t = tokens.shape
        wb, pb = to_bf16_padded(wte), to_bf16_padded(wpe)
        out = torch.empty(b, t, wte.shape[1], dtype=torch.bfloat16, device=wte.device)
        K.embedding_fwd(tokens.contiguous(), wb, pb, out)
        ctx.save_for_backward(tokens)
        ctx.meta = (wte.shape, wpe.shape, wte.dtype)
        return out

    @staticmethod
    def backward(ctx, dout):
        (tokens,) = ctx.saved_tensors
        ws, ps, dt = ctx.meta
        dwte = torch.zeros(ws, dtype=torch.float32, device=dout.device)
        dwpe = torch.zeros(ps, dtype=torch.float32, device=dout.device)
        K.embedding_bwd(tokens.contiguous(), dout.to(torch.bfloat16).contiguous(), dwte, dwpe)
        return None, dwte.to(dt), dwpe.to(dt)


def embedding(tokens: torch.Tensor, wte: torch.Tensor, wpe: torch.Tensor) -> torch.Tensor:
    """``wte[tokens] + wpe[positions]`` in bf16."""
    if wte.is_cuda:
        return _EmbeddingCUDA.apply(tokens, wte, wpe)
    t = tokens.shape[1]
    return (wte.to(torch.bfloat16).float()[tokens] + wpe.to(torch.bfloat16).float()[:t]).to(torch.bfloat16)


class _TiedHeadCUDA(torch.autograd.Function):
    """logits = x @ wte^T (weight tying with the token embedding)."""

    @staticmethod
    def forward(ctx, x, wte):
        lead, c = x.shape[:-1], x.shape[-1]
        x2 = to_bf16_padded(x.reshape(-1, c))
        wb = to_bf16_padded(wte)
        v = wte.shape[0]
        out = torch.empty(x2.shape[0], _pad8(v), dtype=torch.float32, device=x.device)[:, :v]
        K.gemm(x2, wb, out)             # B stored [N=vocab][K=C]: K-major
        ctx.save_for_backward(x2, wb)
        ctx.meta = (lead, c, x.dtype, wte.dtype)
        return out.reshape(*lead, v)

    @staticmethod
    def backward(ctx, dy):
        x2, wb = ctx.saved_tensors
        lead, c, xdt, wdt = ctx.meta
        v = wb.shape[0]
        dy2 = to_bf16_padded(dy.reshape(-1, v))
        dx = torch.empty(x2.shape[0], c, dtype=torch.bfloat16, device=dy.device)
        K.gemm(dy2, wb, dx, b_mn=True)                       # dx = dy @ wte
        dw = torch.zeros(v, c, dtype=torch.float32, device=dy.device)
        K.gemm(dy2, x2, dw, a_mn=True, b_mn=True)            # dwte = dy^T @ x
        return dx.reshape(*lead, c).to(xdt), dw.to(wdt)


def tied_lm_head(x: torch.Tensor, wte: torch.Tensor) -> torch.Tensor:
    if x.is_cuda:
        return _TiedHeadCUDA.apply(x, wte)
    return x.to(torch.bfloat16).float() @ wte.to(torch.bfloat16).float().t()

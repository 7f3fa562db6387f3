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

"""Thin tensor-level wrappers over the C ABI (no autograd here; see ``ops/functional.py``).

Every function launches on the current PyTorch CUDA stream, so the calls compose with
``torch.cuda.graph`` capture and with side streams.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import _lib
from ._lib import AdamWHParams, GemmArgs

ACT_NONE, ACT_SILU, ACT_GELU = 0, 1, 2
_ACT = {None: 0, "none": 0, "silu": 1, "gelu": 2, 0: 0, 1: 1, 2: 2}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _ld(t: torch.Tensor) -> int:
    assert t.dim() == 2 and t.stride(1) == 1, f"need a row-major 2-D view, got {t.shape} {t.stride()}"
    return t.stride(0) if t.shape[0] > 1 else max(t.stride(0), t.shape[1])


def gemm(
    a: torch.Tensor,
    b: torch.Tensor | Sequence[torch.Tensor] | Sequence[int],
    out: torch.Tensor,
    *,
    a_mn: bool = False,
    b_mn: bool = False,
    accumulate: bool = False,
    alpha: float = 1.0,
    bias: Optional[torch.Tensor] = None,
    act=None,
    pre: Optional[torch.Tensor] = None,
    resid: Optional[torch.Tensor] = None,
    bwd: bool = False,
    aux: Optional[torch.Tensor] = None,
    drop_key: Optional[torch.Tensor] = None,
    drop_salt: int = 0,
    drop_rate: float = 0.0,
    drop_salts: Optional[Sequence[int]] = None,
    drop_rows_per_salt: int = 0,
    shard_mode: int = 0,
    shard_len: int = 0,
    rot: int = 0,
    b_shape: Optional[tuple[int, int]] = None,
    ldb: Optional[int] = None,
    block_n: int = 0,
    pair_mode: int = 0,
    split_k: int = 0,
    dbias: Optional[torch.Tensor] = None,
    xent_labels: Optional[torch.Tensor] = None,
    xent_metrics: Optional[torch.Tensor] = None,
    xent_scale: float = 1.0,
) -> torch.Tensor:
    """``out (+)= epilogue(A @ B)`` on tcgen05 (see ``include/b200dist.h`` for the layouts).

    ``a``: [M,K] (``a_mn=False``) or [K,M] (``a_mn=True``); ``b``: [N,K] (``b_mn=False``) or
    [K,N] (``b_mn=True``). ``b`` may be a list of shard tensors or raw device pointers (peer
    memory); then ``b_shape`` is the logical (rows, cols) of the *whole* B in its storage
    orientation and ``ldb`` its row pitch.

    ``xent_labels`` (int64 [M]) turns a classifier head with N <= 32 classes into head + softmax
    cross-entropy in one launch: ``out`` (bf16) receives ``dlogits`` scaled by ``xent_scale``,
    ``xent_metrics[0:2]`` accumulates (loss sum, #correct), ``dbias`` the column sums of ``dlogits``.
    """
    args = GemmArgs()
    assert a.dtype == torch.bfloat16 and a.is_cuda
    if a_mn:
        k, m = a.shape
    else:
        m, k = a.shape
    if isinstance(b, torch.Tensor):
        assert b.dtype == torch.bfloat16
        srcs = [b.data_ptr()]
        rows, cols = b.shape
        ldb_ = _ld(b)
    else:
        srcs = [s.data_ptr() if isinstance(s, torch.Tensor) else int(s) for s in b]
        assert b_shape is not None and ldb is not None
        rows, cols = b_shape
        ldb_ = ldb
    if b_mn:
        kb, n = rows, cols
    else:
        n, kb = rows, cols
    assert kb == k, f"inner dimensions differ: A has K={k}, B has K={kb}"
    assert out.shape[0] == m and out.shape[1] == n, f"out is {tuple(out.shape)}, want {(m, n)}"
    assert out.dtype in (torch.float32, torch.bfloat16)
    args.a = a.data_ptr()
    for i, s in enumerate(srcs):
        args.b_src[i] = s
    args.d = out.data_ptr()
    args.m, args.n, args.k = m, n, k
    args.lda, args.ldb, args.ldd = _ld(a), ldb_, _ld(out)
    args.a_mn, args.b_mn = int(a_mn), int(b_mn)
    args.b_nsrc, args.b_shard_mode, args.b_shard_len, args.b_rot = len(srcs), shard_mode, shard_len, rot
    args.d_fp32 = int(out.dtype == torch.float32)
    args.accumulate = int(accumulate)
    args.alpha = alpha
    args.bias = _ptr(bias)
    if pre is not None:
        args.pre, args.ldpre = pre.data_ptr(), _ld(pre)
    args.act = _ACT[act]
    if resid is not None:
        args.resid, args.ldres = resid.data_ptr(), _ld(resid)
    args.bwd = int(bwd)
    if aux is not None:
        args.aux, args.ldaux = aux.data_ptr(), _ld(aux)
    if drop_key is not None and drop_rate > 0.0:
        args.drop_key = drop_key.data_ptr()
        args.drop_salt = drop_salt & 0xFFFFFFFFFFFFFFFF
        args.drop_rate = drop_rate
        if drop_salts is not None:
            assert len(drop_salts) <= 16 and drop_rows_per_salt > 0
            args.drop_nsalt, args.drop_rows_per_salt = len(drop_salts), drop_rows_per_salt
            for i, sv in enumerate(drop_salts):
                args.drop_salts[i] = sv & 0xFFFFFFFFFFFFFFFF
    args.block_n = block_n
    args.pair_mode = pair_mode
    args.split_k = split_k
    if dbias is not None:
        assert dbias.dtype == torch.float32 and dbias.numel() >= n
        args.dbias = dbias.data_ptr()
    if xent_labels is not None:
        assert xent_labels.dtype == torch.int64 and xent_labels.numel() >= m and out.dtype == torch.bfloat16
        args.xent_labels = xent_labels.data_ptr()
        args.xent_metrics = _ptr(xent_metrics)
        args.xent_scale = float(xent_scale)
    _lib.check(_lib.lib().b200_gemm_bf16(C.byref(args), _stream()))
    return out


def softmax_xent(logits: torch.Tensor, labels: torch.Tensor, *, grad_scale: float,
                 dlogits: Optional[torch.Tensor], metrics: Optional[torch.Tensor],
                 row_loss: Optional[torch.Tensor] = None, dbias: Optional[torch.Tensor] = None) -> None:
    """fp32 logits [rows, classes] -> dlogits bf16, metrics[0] += loss sum, metrics[1] += #correct;
    ``dbias[c] += sum_rows dlogits[:, c]`` when given."""
    assert logits.dtype == torch.float32 and labels.dtype == torch.int64
    rows, classes = logits.shape
    _lib.check(_lib.lib().b200_softmax_xent_dbias(
        logits.data_ptr(), _ld(logits), labels.data_ptr(), rows, classes, grad_scale,
        _ptr(dlogits), _ld(dlogits) if dlogits is not None else 0, _ptr(metrics), _ptr(row_loss), _ptr(dbias),
        _stream()))


def softmax_xent_bf16_(logits: torch.Tensor, labels: torch.Tensor, *, grad_scale: float,
                       metrics: Optional[torch.Tensor], row_loss: Optional[torch.Tensor] = None) -> None:
    """bf16 logits are replaced in place by their gradient."""
    assert logits.dtype == torch.bfloat16 and labels.dtype == torch.int64
    rows, classes = logits.shape
    _lib.check(_lib.lib().b200_softmax_xent_bf16_inplace(
        logits.data_ptr(), _ld(logits), labels.data_ptr(), rows, classes, grad_scale,
        _ptr(metrics), _ptr(row_loss), _stream()))


def colsum(x: torch.Tensor, out: torch.Tensor, accumulate: bool = False) -> torch.Tensor:
    rows, cols = x.shape
    assert x.dtype == torch.bfloat16 and out.dtype == torch.float32 and out.numel() >= cols
    _lib.check(_lib.lib().b200_colsum_bf16(x.data_ptr(), _ld(x), rows, cols, out.data_ptr(),
                                            int(accumulate), _stream()))
    return out


def adamw_(p: torch.Tensor, m: torch.Tensor, v: torch.Tensor, g: torch.Tensor,
           p_bf16: Optional[torch.Tensor], step_count: torch.Tensor, *, lr: float, beta1: float = 0.9,
           beta2: float = 0.999, eps: float = 1e-8, weight_decay: float = 1e-4, grad_scale: float = 1.0,
           bump_step: bool = True, zero_grad: bool = False) -> None:
    """In-place AdamW over flat fp32 buffers (optax.adamw semantics), optional bf16 copy."""
    hp = AdamWHParams(lr, beta1, beta2, eps, weight_decay, grad_scale)
    assert step_count.dtype == torch.int32
    _lib.check(_lib.lib().b200_adamw(p.data_ptr(), m.data_ptr(), v.data_ptr(), g.data_ptr(),
                                      _ptr(p_bf16), p.numel(), C.byref(hp), step_count.data_ptr(),
                                      int(bump_step), int(zero_grad), _stream()))


def cast_f32_bf16(src: torch.Tensor, dst: torch.Tensor) -> torch.Tensor:
    _lib.check(_lib.lib().b200_cast_f32_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _stream()))
    return dst


def cast_bf16_f32(src: torch.Tensor, dst: torch.Tensor) -> torch.Tensor:
    _lib.check(_lib.lib().b200_cast_bf16_f32(src.data_ptr(), dst.data_ptr(), src.numel(), _stream()))
    return dst


def scale_(x: torch.Tensor, s: float) -> torch.Tensor:
    _lib.check(_lib.lib().b200_scale_f32(x.data_ptr(), x.numel(), s, _stream()))
    return x


def rng_split_(key: torch.Tensor, out_step: Optional[torch.Tensor], salt_next: int, salt_step: int) -> None:
    _lib.check(_lib.lib().b200_rng_split(key.data_ptr(), _ptr(out_step), salt_next & (2**64 - 1),
                                          salt_step & (2**64 - 1), _stream()))


def rng_split_bump_(key: torch.Tensor, salt_next: int, step_count: torch.Tensor) -> None:
    """``key <- mix64(key ^ salt_next)`` and ``step_count += 1`` in one launch."""
    _lib.check(_lib.lib().b200_rng_split_bump(key.data_ptr(), None, salt_next & (2**64 - 1), 0, step_count.data_ptr(),
                                               _stream()))


def cast_zero(src: Optional[torch.Tensor], dst: Optional[torch.Tensor], z0: Optional[torch.Tensor] = None,
              z1: Optional[torch.Tensor] = None, *, wait_signal: Optional[int] = None, peer_signals=None,
              rank: int = 0, world: int = 1) -> None:
    """bf16 cast of ``src`` into ``dst`` plus zeroing of up to two fp32 regions, one launch.
    ``wait_signal`` (address of this rank's signal buffer) makes the kernel first wait until every
    peer has published its part of the previous fused update (csrc/comm.cuh)."""
    _lib.check(_lib.lib().b200_cast_zero_wait(_ptr(src), _ptr(dst), src.numel() if src is not None else 0,
                                               _ptr(z0), z0.numel() if z0 is not None else 0,
                                               _ptr(z1), z1.numel() if z1 is not None else 0,
                                               wait_signal if world > 1 else None, peer_signals, rank, world, _stream()))


def act_dropout_fwd(x: torch.Tensor, y: torch.Tensor, act, drop_key, drop_salt: int, drop_rate: float):
    rows, cols = x.shape
    _lib.check(_lib.lib().b200_act_dropout_fwd(x.data_ptr(), y.data_ptr(), rows, cols, _ld(x), _ACT[act],
                                                _ptr(drop_key), drop_salt & (2**64 - 1), drop_rate, _stream()))
    return y


def act_dropout_bwd(x: torch.Tensor, dy: torch.Tensor, dx: torch.Tensor, act, drop_key, drop_salt: int,
                    drop_rate: float):
    rows, cols = x.shape
    _lib.check(_lib.lib().b200_act_dropout_bwd(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), rows, cols, _ld(x),
                                                _ACT[act], _ptr(drop_key), drop_salt & (2**64 - 1), drop_rate,
                                                _stream()))
    return dx


def layernorm_fwd(x, gamma, beta, y, mean, rstd, eps: float = 1e-5):
    rows, cols = x.shape
    _lib.check(_lib.lib().b200_layernorm_fwd(x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), y.data_ptr(),
                                              mean.data_ptr(), rstd.data_ptr(), rows, cols, eps, _stream()))
    return y


def layernorm_bwd(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, dresid=None, dxsum=None):
    rows, cols = x.shape
    _lib.check(_lib.lib().b200_layernorm_bwd_dxsum(dy.data_ptr(), x.data_ptr(), gamma.data_ptr(), mean.data_ptr(),
                                                    rstd.data_ptr(), _ptr(dresid), dx.data_ptr(), dgamma.data_ptr(),
                                                    dbeta.data_ptr(), _ptr(dxsum), rows, cols, _stream()))
    return dx


def embedding_fwd(tokens, wte, wpe, out):
    b, t = tokens.shape
    _lib.check(_lib.lib().b200_embedding_fwd(tokens.data_ptr(), wte.data_ptr(), wpe.data_ptr(), out.data_ptr(),
                                              b, t, wte.shape[1], _stream()))
    return out


def embedding_bwd(tokens, dout, dwte, dwpe):
    b, t = tokens.shape
    _lib.check(_lib.lib().b200_embedding_bwd(tokens.data_ptr(), dout.data_ptr(), dwte.data_ptr(),
                                              dwpe.data_ptr(), b, t, dout.shape[-1], _stream()))


def attention_fwd(qkv, out, lse, batch: int, seq: int, heads: int):
    _lib.check(_lib.lib().b200_attention_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), batch, seq, heads,
                                              _stream()))
    return out


def attention_bwd_scratch_numel(batch: int, seq: int, heads: int) -> int:
    """fp32 elements of scratch that ``attention_bwd`` needs."""
    return int(_lib.lib().b200_attention_bwd_scratch_bytes(batch, seq, heads)) // 4


def attention_bwd(qkv, out, dout, lse, dqkv, scratch, batch: int, seq: int, heads: int):
    assert scratch.dtype == torch.float32 and scratch.numel() >= attention_bwd_scratch_numel(batch, seq, heads)
    _lib.check(_lib.lib().b200_attention_bwd(qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(),
                                              dqkv.data_ptr(), scratch.data_ptr(), batch, seq, heads, _stream()))
    return dqkv


def embedding_fwd_sharded(tokens, wte_peers, wpe_peers, out, dim: int):
    """Embedding lookup over row-sharded tables that live in peer memory (FSDP): a token's row is
    read from the rank that owns it. ``*_peers`` = (list of device pointers, rows per shard)."""
    b, t = tokens.shape
    (wp, wrows), (pp, prows) = wte_peers, wpe_peers
    wa = (C.c_void_p * 8)(*(list(wp) + [None] * (8 - len(wp))))
    pa = (C.c_void_p * 8)(*(list(pp) + [None] * (8 - len(pp))))
    _lib.check(_lib.lib().b200_embedding_fwd_sharded(tokens.data_ptr(), wa, len(wp), wrows, pa, len(pp), prows,
                                                      out.data_ptr(), b, t, dim, _stream()))
    return out

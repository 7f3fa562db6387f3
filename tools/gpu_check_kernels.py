"""GPU numerics + timing sweep for the raw kernels, against plain PyTorch fp32 references.

Run on a B200:  python tools/gpu_check_kernels.py            (all groups, each in a subprocess)
                python tools/gpu_check_kernels.py --group gemm_kmajor
Results go to stdout and gpurun_out/kernels_check.log. Each group runs in its own process with
a timeout so that a faulting kernel does not hide the other groups.
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

GROUPS = ["gemm_kmajor", "gemm_mnmajor", "gemm_epilogue", "gemm_xent", "gemm_sharded", "elementwise", "transformer",
          "attention", "gemm_perf", "gemm_epi_perf"]


class _Err(float):
    """Max-norm error that also carries the element-wise excess (see ``_rel_err``)."""
    elem = 0.0
    nbad_at = None


def _rel_err(got, ref):
    """Largest error relative to the largest reference magnitude. The returned float also carries
    ``elem`` = max over elements of (|got - ref| - 0.1 |ref|) / rms(ref), which ``_report`` holds against
    25 x the check's tolerance: the max-norm alone lets a wrong value hide in a small-magnitude region (a
    flipped sign there is far below the global scale), the element-wise bound scaled by the same tolerance
    does not."""
    import torch
    got = got.float()
    ref = ref.float()
    denom = ref.abs().max().clamp_min(1e-6)
    d = (got - ref).abs()
    err = _Err((d.max() / denom).item())
    rms = ref.pow(2).mean().sqrt().clamp_min(1e-12)
    err.elem = ((d - 0.10 * ref.abs()).clamp_min(0) / rms).max().item() if d.numel() else 0.0
    return err


def _report(name, err, tol):
    elem = getattr(err, "elem", 0.0)
    elem_ok = elem == elem and elem <= 25.0 * tol
    ok = err == err and err <= tol and elem_ok
    extra = "" if elem_ok else f" element-wise excess {elem:.3e} rms (bound {25.0 * tol:.1e})"
    print(f"  [{'ok' if ok else 'FAIL'}] {name}: rel_err={float(err):.3e} (tol {tol:.1e}){extra}", flush=True)
    return ok


def _time(fn, iters=20, warm=3):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def group_gemm_kmajor():
    import torch
    from jax_distributed_tuts_b200.ops import kernels as K
    ok = True
    torch.manual_seed(0)
    for (m, n, k, bn) in [(128, 64, 64, 64), (128, 128, 64, 128), (128, 256, 64, 256), (128, 128, 256, 128),
                          (256, 256, 512, 0), (384, 320, 200, 0), (4, 16, 784, 0), (1000, 1000, 1000, 0),
                          (4096, 4096, 1024, 256), (4096, 4096, 1024, 128)]:
        kp = (k + 7) // 8 * 8
        a = torch.randn(m, kp, device="cuda", dtype=torch.bfloat16)[:, :k]
        b = torch.randn(n, kp, device="cuda", dtype=torch.bfloat16)[:, :k]
        out = torch.full((m, n), float("nan"), device="cuda", dtype=torch.float32)
        K.gemm(a, b, out, block_n=bn)
        ref = a.float() @ b.float().t()
        ok &= _report(f"TN fp32 out m={m} n={n} k={k} bn={bn}", _rel_err(out, ref), 2e-3)
        if bn != 64 and n > 64:
            for pm in (1, 2):
                out.fill_(float("nan"))
                K.gemm(a, b, out, block_n=bn, pair_mode=pm)
                ok &= _report(f"   pair_mode={pm}", _rel_err(out, ref), 2e-3)
        if not ok and m * n <= 128 * 256:
            bad = ((out - ref).abs() > 1e-2 * ref.abs().max()).nonzero()
            print("   first bad idx:", bad[:8].tolist(), "n_bad", bad.shape[0])
    # bf16 out + accumulate
    a = torch.randn(512, 256, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(384, 256, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(512, 384, device="cuda", dtype=torch.bfloat16)
    K.gemm(a, b, out)
    ok &= _report("TN bf16 out", _rel_err(out, a.float() @ b.float().t()), 1e-2)
    acc = torch.ones(512, 384, device="cuda", dtype=torch.float32)
    K.gemm(a, b, acc, accumulate=True)
    K.gemm(a, b, acc, accumulate=True, alpha=0.5)
    ok &= _report("TN accumulate", _rel_err(acc, 1.0 + 1.5 * (a.float() @ b.float().t())), 2e-3)
    return ok


def group_gemm_mnmajor():
    import torch
    from jax_distributed_tuts_b200.ops import kernels as K
    ok = True
    torch.manual_seed(1)
    for (m, n, k) in [(128, 64, 64), (128, 128, 64), (128, 256, 128), (256, 512, 320), (300, 200, 136),
                      (16, 512, 784), (784, 512, 16), (512, 16, 16), (2048, 3072, 768)]:
        r8 = lambda v: (v + 7) // 8 * 8
        a = torch.randn(m, r8(k), device="cuda", dtype=torch.bfloat16)[:, :k]
        at = torch.zeros(k, r8(m), device="cuda", dtype=torch.bfloat16)[:, :m]   # [K, M] storage
        at.copy_(a.t())
        b = torch.randn(k, r8(n), device="cuda", dtype=torch.bfloat16)[:, :n]   # [K, N] storage
        bt = torch.zeros(n, r8(k), device="cuda", dtype=torch.bfloat16)[:, :k]   # [N, K]
        bt.copy_(b.t())
        ref = a.float() @ b.float()
        out = torch.full((m, n), float("nan"), device="cuda", dtype=torch.float32)
        K.gemm(a, b, out, b_mn=True)
        ok &= _report(f"NN (B MN-major) m={m} n={n} k={k}", _rel_err(out, ref), 2e-3)
        out.fill_(float("nan"))
        K.gemm(at, bt, out, a_mn=True)
        ok &= _report(f"A MN-major, B K-major m={m} n={n} k={k}", _rel_err(out, ref), 2e-3)
        out.fill_(float("nan"))
        K.gemm(at, b, out, a_mn=True, b_mn=True)
        ok &= _report(f"both MN-major (wgrad) m={m} n={n} k={k}", _rel_err(out, ref), 2e-3)
    # split-K (small M x N grid, long K): automatic and forced, with and without accumulation
    m, n, k = 768, 1024, 4096
    at = torch.randn(k, m, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(k, n, device="cuda", dtype=torch.bfloat16)
    ref = at.float().t() @ b.float()
    for sk in (0, 1, 3, 8):
        for pm in (1, 2):
            acc = torch.full((m, n), float("nan"), device="cuda")
            K.gemm(at, b, acc, a_mn=True, b_mn=True, split_k=sk, pair_mode=pm)
            ok &= _report(f"wgrad split_k={sk} pair_mode={pm}", _rel_err(acc, ref), 2e-3)
            K.gemm(at, b, acc, a_mn=True, b_mn=True, split_k=sk, pair_mode=pm, accumulate=True)
            ok &= _report(f"wgrad split_k={sk} pair_mode={pm} accumulate", _rel_err(acc, 2 * ref), 2e-3)
    # padded pitch (N=10 stored with ld 16)
    x = torch.randn(64, 512, device="cuda", dtype=torch.bfloat16)
    wbuf = torch.zeros(512, 16, device="cuda", dtype=torch.bfloat16)
    w = wbuf[:, :10]
    w.copy_(torch.randn(512, 10))
    out = torch.empty(64, 10, device="cuda", dtype=torch.float32)
    K.gemm(x, w, out, b_mn=True)
    ok &= _report("N=10 padded pitch fwd", _rel_err(out, x.float() @ w.float()), 2e-3)
    dybuf = torch.zeros(64, 16, device="cuda", dtype=torch.bfloat16)
    dy = dybuf[:, :10]
    dy.copy_(torch.randn(64, 10))
    dwbuf = torch.zeros(512, 16, device="cuda", dtype=torch.float32)
    dw = dwbuf[:, :10]
    K.gemm(x, dy, dw, a_mn=True, b_mn=True, accumulate=True)
    ok &= _report("N=10 padded pitch wgrad", _rel_err(dw, x.float().t() @ dy.float()), 2e-3)
    dx = torch.empty(64, 512, device="cuda", dtype=torch.bfloat16)
    K.gemm(dy, w, dx)      # dx = dy @ w^T : B stored [N=512][K=10]
    ok &= _report("K=10 padded pitch dgrad", _rel_err(dx, dy.float() @ w.float().t()), 1e-2)
    return ok


def group_gemm_epilogue():
    import torch
    import torch.nn.functional as F
    from jax_distributed_tuts_b200.ops import kernels as K
    ok = True
    torch.manual_seed(2)
    m, n, k = 256, 512, 784
    x = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
    w = (torch.randn(k, n, device="cuda") / k ** 0.5).bfloat16()
    bias = torch.randn(n, device="cuda").bfloat16()
    pre = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    for act, f in (("silu", F.silu), ("gelu", lambda t: F.gelu(t, approximate="tanh")), (None, lambda t: t)):
        K.gemm(x, w, out, b_mn=True, bias=bias, act=act, pre=pre)
        ref_pre = x.float() @ w.float() + bias.float()
        ok &= _report(f"bias+{act} pre", _rel_err(pre, ref_pre), 1e-2)
        ok &= _report(f"bias+{act} out", _rel_err(out, f(ref_pre)), 1.5e-2)
    # residual
    res = torch.randn(m, n, device="cuda", dtype=torch.bfloat16)
    K.gemm(x, w, out, b_mn=True, bias=bias, resid=res)
    ok &= _report("bias+residual", _rel_err(out, x.float() @ w.float() + bias.float() + res.float()), 1e-2)
    # dropout fwd: fraction kept and scaling; bwd regenerates the same mask
    key = torch.tensor([1234567], device="cuda", dtype=torch.int64)
    K.gemm(x, w, out, b_mn=True, bias=bias, act="silu", pre=pre, drop_key=key, drop_salt=7, drop_rate=0.1)
    ref = F.silu(pre.float())
    kept = out != 0
    frac = kept.float().mean().item()
    print(f"  dropout keep fraction {frac:.4f} (want ~0.9)")
    ok &= abs(frac - 0.9) < 0.01
    ok &= _report("dropout kept values scaled", _rel_err(out[kept], (ref / 0.9)[kept]), 1.5e-2)
    dy = torch.randn(m, n, device="cuda", dtype=torch.bfloat16)
    # bwd epilogue: dpre = (dy_in) * mask/keep * silu'(pre); use identity GEMM input to isolate epilogue
    eye = torch.eye(n, device="cuda", dtype=torch.bfloat16)
    dpre = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    K.gemm(dy, eye, dpre, bwd=True, aux=pre, act="silu", drop_key=key, drop_salt=7, drop_rate=0.1)
    p32 = pre.float()
    s = torch.sigmoid(p32)
    ref_d = dy.float() * (kept.float() / 0.9) * (s * (1 + p32 * (1 - s)))
    ok &= _report("bwd epilogue silu'+dropout mask", _rel_err(dpre, ref_d), 1.5e-2)
    K.gemm(dy, eye, dpre, bwd=True, aux=pre, act="gelu")
    pg = pre.float().requires_grad_(True)
    F.gelu(pg, approximate="tanh").backward(dy.float())
    ok &= _report("bwd epilogue gelu'", _rel_err(dpre, pg.grad), 1.5e-2)
    # dbias: column sums of the stored output from inside the epilogue (fast and generic variants)
    for (mm, nn, kk, kw) in [(64, 512, 16, dict()), (64, 512, 16, dict(bwd=True, act="silu")),
                             (300, 768, 256, dict(bwd=True, act="gelu")), (1024, 3072, 768, dict(bwd=True, act="gelu"))]:
        a2 = torch.randn(mm, kk, device="cuda", dtype=torch.bfloat16)
        b2 = torch.randn(nn, kk, device="cuda", dtype=torch.bfloat16) * 0.1
        aux2 = torch.randn(mm, nn, device="cuda", dtype=torch.bfloat16)
        o2 = torch.empty(mm, nn, device="cuda", dtype=torch.bfloat16)
        for drop in (0.0, 0.1):
            db = torch.zeros(nn, device="cuda")
            K.gemm(a2, b2, o2, dbias=db, drop_key=key, drop_rate=drop, drop_salts=[3, 4], drop_rows_per_salt=max(1, mm // 2),
                   **kw, **(dict(aux=aux2) if kw.get("bwd") else {}))
            ok &= _report(f"dbias m={mm} n={nn} {kw.get('act')} drop={drop}", _rel_err(db, o2.float().sum(0)), 1e-5)
    # different salt -> different mask
    out2 = torch.empty_like(out)
    K.gemm(x, w, out2, b_mn=True, bias=bias, act="silu", drop_key=key, drop_salt=8, drop_rate=0.1)
    same = ((out2 != 0) == kept).float().mean().item()
    print(f"  masks with different salt agree on {same:.3f} of elements (want ~0.82)")
    ok &= same < 0.9
    return ok


def group_gemm_xent():
    """Classifier head + softmax cross-entropy in one GEMM launch vs. PyTorch fp32."""
    import torch
    import torch.nn.functional as F
    from jax_distributed_tuts_b200.ops import kernels as K
    ok = True
    torch.manual_seed(9)
    for (m, k, n) in [(16, 512, 10), (200, 784, 32), (130, 64, 3)]:
        x = torch.randn(m, k, device="cuda").bfloat16()
        ld = (n + 7) // 8 * 8
        w = (torch.randn(k, ld, device="cuda") / k ** 0.5).bfloat16()[:, :n]
        bias = torch.randn(n, device="cuda").bfloat16()
        labels = torch.randint(0, n, (m,), device="cuda")
        dl = torch.zeros(m, ld, device="cuda", dtype=torch.bfloat16)[:, :n]
        metrics = torch.zeros(2, device="cuda")
        db = torch.zeros(64, device="cuda")
        scale = 1.0 / m
        K.gemm(x, w, dl, b_mn=True, bias=bias, xent_labels=labels, xent_metrics=metrics, xent_scale=scale, dbias=db)
        logits = (x.float() @ w.float() + bias.float()).requires_grad_(True)
        loss = F.cross_entropy(logits, labels, reduction="sum")
        (loss * scale).backward()
        ok &= _report(f"xent head {m}x{k}x{n} loss sum", abs(metrics[0].item() - loss.item()) / max(1.0, loss.item()), 2e-3)
        ok &= _report("xent head correct", abs(metrics[1].item() - (logits.argmax(-1) == labels).sum().item()), 1.0)
        ok &= _report("xent head dlogits", _rel_err(dl, logits.grad), 1.5e-2)
        ok &= _report("xent head dbias", _rel_err(db[:n], logits.grad.sum(0)), 1.5e-2)
    return ok


def group_gemm_sharded():
    import torch
    from jax_distributed_tuts_b200.ops import kernels as K
    ok = True
    torch.manual_seed(3)
    # K-sharded MN-major B (Flax (in,out) kernel sharded on dim 0), shard_len not a multiple of 64
    for (m, n, k, p) in [(64, 512, 784, 8), (256, 256, 1024, 4), (128, 3072, 768, 8)]:
        x = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
        w = (torch.randn(k, n, device="cuda") / k ** 0.5).bfloat16()
        shards = [s.contiguous() for s in w.chunk(p, dim=0)]
        out = torch.full((m, n), float("nan"), device="cuda", dtype=torch.float32)
        for rot in (0, p - 1):
            K.gemm(x, shards, out, b_mn=True, shard_mode=1, shard_len=k // p, rot=rot,
                   b_shape=(k, n), ldb=n)
            torch.cuda.synchronize()
            ok &= _report(f"K-sharded fwd m={m} n={n} k={k} p={p} rot={rot}", _rel_err(out, x.float() @ w.float()), 2e-3)
        # N-sharded MN-major B (kernel sharded on dim 1)
        shards_n = [s.contiguous() for s in w.chunk(p, dim=1)]
        out.fill_(float("nan"))
        K.gemm(x, shards_n, out, b_mn=True, shard_mode=2, shard_len=n // p, rot=1,
               b_shape=(k, n), ldb=n // p)
        ok &= _report(f"N-sharded fwd m={m} n={n} k={k} p={p}", _rel_err(out, x.float() @ w.float()), 2e-3)
        # dgrad with the same shards: dx = dy @ w^T ; B stored [N'=k][K'=n] K-major
        dy = torch.randn(m, n, device="cuda", dtype=torch.bfloat16)
        dx = torch.full((m, k), float("nan"), device="cuda", dtype=torch.float32)
        K.gemm(dy, shards, dx, shard_mode=2, shard_len=k // p, rot=2 % p, b_shape=(k, n), ldb=n)
        ok &= _report(f"dgrad, dim-0 shards m={m}", _rel_err(dx, dy.float() @ w.float().t()), 2e-3)
        dx.fill_(float("nan"))
        K.gemm(dy, shards_n, dx, shard_mode=1, shard_len=n // p, rot=1, b_shape=(k, n), ldb=n // p)
        ok &= _report(f"dgrad, dim-1 shards m={m}", _rel_err(dx, dy.float() @ w.float().t()), 2e-3)
    return ok


def group_elementwise():
    import torch
    import torch.nn.functional as F
    from jax_distributed_tuts_b200.ops import kernels as K
    ok = True
    torch.manual_seed(4)
    for rows, classes, ld in [(4, 10, 10), (128, 10, 16), (1000, 1000, 1000), (77, 3333, 3336)]:
        buf = torch.randn(rows, ld, device="cuda") * 3
        logits = buf[:, :classes]
        labels = torch.randint(0, classes, (rows,), device="cuda")
        dl_buf = torch.zeros(rows, (classes + 7) // 8 * 8, device="cuda", dtype=torch.bfloat16)
        dl = dl_buf[:, :classes]
        metrics = torch.zeros(2, device="cuda")
        rl = torch.empty(rows, device="cuda")
        K.softmax_xent(logits, labels, grad_scale=1.0 / rows, dlogits=dl, metrics=metrics, row_loss=rl)
        lg = logits.clone().requires_grad_(True)
        loss = F.cross_entropy(lg, labels, reduction="none")
        loss.mean().backward()
        ok &= _report(f"xent loss rows={rows} C={classes}", _rel_err(rl, loss.detach()), 1e-4)
        ok &= _report("xent loss sum", abs(metrics[0].item() - loss.sum().item()) / max(1.0, loss.sum().item()), 1e-4)
        ok &= _report("xent correct", abs(metrics[1].item() - (lg.argmax(-1) == labels).sum().item()), 0.0)
        ok &= _report("xent dlogits", _rel_err(dl, lg.grad), 1e-2)
    # bf16 in-place vocabulary variant
    # bf16 in-place variant: 50257 takes the register-resident kernel, the other two the generic block kernel
    for rows, classes in [(300, 50257), (64, 1000), (48, 60000)]:
        ld = (classes + 7) // 8 * 8
        buf = (torch.randn(rows, ld, device="cuda") * 2).bfloat16()
        logits = buf[:, :classes]
        ref_in = logits.float().clone().requires_grad_(True)
        labels = torch.randint(0, classes, (rows,), device="cuda")
        labels[0] = classes - 1           # label in the scalar tail
        loss = F.cross_entropy(ref_in, labels, reduction="none")
        loss.mean().backward()
        metrics = torch.zeros(2, device="cuda")
        rl = torch.empty(rows, device="cuda")
        K.softmax_xent_bf16_(logits, labels, grad_scale=1.0 / rows, metrics=metrics, row_loss=rl)
        ok &= _report(f"xent bf16 C={classes} loss", _rel_err(rl, loss.detach()), 1e-3)
        ok &= _report("xent bf16 dlogits", _rel_err(logits, ref_in.grad), 1e-2)
        ok &= _report("xent bf16 correct", abs(metrics[1].item() - (ref_in.argmax(-1) == labels).sum().item()), 0.0)
    # colsum
    for rows, cols in [(4, 10), (128, 512), (8192, 768), (1000, 50)]:
        x = torch.randn(rows, (cols + 7) // 8 * 8, device="cuda", dtype=torch.bfloat16)[:, :cols]
        out = torch.full((cols,), 1.0, device="cuda")
        K.colsum(x, out, accumulate=False)
        ok &= _report(f"colsum {rows}x{cols}", _rel_err(out, x.float().sum(0)), 1e-3)
        K.colsum(x, out, accumulate=True)
        ok &= _report(f"colsum accumulate {rows}x{cols}", _rel_err(out, 2 * x.float().sum(0)), 1e-3)
    # adamw vs torch reference of the optax formula
    n = 407050 + 6
    p = torch.randn(n, device="cuda"); g = torch.randn(n, device="cuda") * 0.1
    m = torch.zeros(n, device="cuda"); v = torch.zeros(n, device="cuda")
    pb = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    step = torch.zeros(1, device="cuda", dtype=torch.int32)
    p_ref, m_ref, v_ref = p.clone().double(), m.clone().double(), v.clone().double()
    lr, b1, b2, eps, wd = 1e-3, 0.9, 0.999, 1e-8, 1e-4
    for t in range(1, 4):
        K.adamw_(p, m, v, g, pb, step, lr=lr, weight_decay=wd, grad_scale=0.25)
        gg = g.double() * 0.25
        m_ref = b1 * m_ref + (1 - b1) * gg
        v_ref = b2 * v_ref + (1 - b2) * gg * gg
        mh = m_ref / (1 - b1 ** t); vh = v_ref / (1 - b2 ** t)
        p_ref = p_ref - lr * (mh / (vh.sqrt() + eps) + wd * p_ref)
    ok &= _report("adamw p (3 steps)", _rel_err(p, p_ref.float()), 1e-5)
    ok &= _report("adamw m", _rel_err(m, m_ref.float()), 1e-5)
    ok &= _report("adamw v", _rel_err(v, v_ref.float()), 1e-4)
    ok &= _report("adamw bf16 copy", _rel_err(pb, p), 4e-3)
    ok &= _report("adamw step counter", abs(step.item() - 3), 0)
    # rng split
    key = torch.tensor([42], device="cuda", dtype=torch.int64)
    stepk = torch.zeros(1, device="cuda", dtype=torch.int64)
    K.rng_split_(key, stepk, 1, 2)
    from jax_distributed_tuts_b200.rng import mix64
    ok &= _report("rng next", float((key.item() & (2**64 - 1)) != mix64(42 ^ 1)), 0)
    ok &= _report("rng step", float((stepk.item() & (2**64 - 1)) != mix64(42 ^ 2)), 0)
    # act/dropout stand-alone
    x = torch.randn(100, 264, device="cuda", dtype=torch.bfloat16)[:, :260]
    y = torch.zeros(100, 264, device="cuda", dtype=torch.bfloat16)[:, :260]
    K.act_dropout_fwd(x, y, "gelu", None, 0, 0.0)
    ok &= _report("gelu fwd", _rel_err(y, F.gelu(x.float(), approximate="tanh")), 1e-2)
    return ok


def group_transformer():
    import torch
    import torch.nn.functional as F
    from jax_distributed_tuts_b200.ops import kernels as K
    ok = True
    torch.manual_seed(5)
    for rows, cols in [(8, 768), (4096, 768), (1000, 1024), (64, 1600)]:
        x = torch.randn(rows, cols, device="cuda").bfloat16()
        gam = (1 + 0.1 * torch.randn(cols, device="cuda")).bfloat16()
        bet = (0.1 * torch.randn(cols, device="cuda")).bfloat16()
        y = torch.empty_like(x); mean = torch.empty(rows, device="cuda"); rstd = torch.empty(rows, device="cuda")
        K.layernorm_fwd(x, gam, bet, y, mean, rstd)
        xr = x.float().requires_grad_(True); gr = gam.float().requires_grad_(True); br = bet.float().requires_grad_(True)
        yr = F.layer_norm(xr, (cols,), gr, br, 1e-5)
        ok &= _report(f"layernorm fwd {rows}x{cols}", _rel_err(y, yr), 1e-2)
        dy = torch.randn(rows, cols, device="cuda").bfloat16()
        yr.backward(dy.float())
        dx = torch.empty_like(x); dg = torch.zeros(cols, device="cuda"); db = torch.zeros(cols, device="cuda")
        K.layernorm_bwd(dy, x, gam, mean, rstd, dx, dg, db)
        ok &= _report("layernorm dx", _rel_err(dx, xr.grad), 1.5e-2)
        ok &= _report("layernorm dgamma", _rel_err(dg, gr.grad), 1e-2)
        ok &= _report("layernorm dbeta", _rel_err(db, br.grad), 1e-2)
    b, t, d, vocab = 4, 128, 768, 50257
    tok = torch.randint(0, vocab, (b, t), device="cuda")
    wte = torch.randn(vocab, d, device="cuda").bfloat16(); wpe = torch.randn(1024, d, device="cuda").bfloat16()
    out = torch.empty(b, t, d, device="cuda", dtype=torch.bfloat16)
    K.embedding_fwd(tok, wte, wpe, out)
    ok &= _report("embedding fwd", _rel_err(out, wte[tok].float() + wpe[:t].float()), 1e-2)
    dout = torch.randn(b, t, d, device="cuda").bfloat16()
    dwte = torch.zeros(vocab, d, device="cuda"); dwpe = torch.zeros(1024, d, device="cuda")
    K.embedding_bwd(tok, dout, dwte, dwpe)
    ref_wte = torch.zeros(vocab, d, device="cuda").index_add_(0, tok.flatten(), dout.float().view(-1, d))
    ok &= _report("embedding dwte", _rel_err(dwte, ref_wte), 1e-4)
    ok &= _report("embedding dwpe", _rel_err(dwpe[:t], dout.float().sum(0)), 1e-4)
    return ok


def group_attention():
    import torch
    import torch.nn.functional as F
    from jax_distributed_tuts_b200.ops import kernels as K
    ok = True
    torch.manual_seed(6)
    for (b, t, h) in [(1, 128, 1), (2, 256, 3), (2, 1024, 12)]:
        qkv = torch.randn(b, t, 3, h, 64, device="cuda").bfloat16()
        out = torch.empty(b, t, h, 64, device="cuda", dtype=torch.bfloat16)
        lse = torch.empty(b, h, t, device="cuda")
        K.attention_fwd(qkv, out, lse, b, t, h)
        q, k, v = [qkv[:, :, i].float().permute(0, 2, 1, 3).detach().requires_grad_(True) for i in range(3)]
        ref = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        ok &= _report(f"attention fwd b={b} t={t} h={h}", _rel_err(out.permute(0, 2, 1, 3), ref), 1.5e-2)
        dout = torch.randn(b, t, h, 64, device="cuda").bfloat16()
        ref.backward(dout.float().permute(0, 2, 1, 3))
        dqkv = torch.empty_like(qkv)
        scratch = torch.empty(K.attention_bwd_scratch_numel(b, t, h), device="cuda")
        K.attention_bwd(qkv, out, dout, lse, dqkv, scratch, b, t, h)
        for i, (nm, gref) in enumerate((("dq", q.grad), ("dk", k.grad), ("dv", v.grad))):
            ok &= _report(f"attention {nm}", _rel_err(dqkv[:, :, i].permute(0, 2, 1, 3), gref), 2e-2)
    # timing at the GPT-2 small bench shape (per layer): 8 and 16 sequences of 1024, 12 heads
    for b in (8, 16):
        t, h = 1024, 12
        qkv = torch.randn(b, t, 3, h, 64, device="cuda").bfloat16()
        out = torch.empty(b, t, h, 64, device="cuda", dtype=torch.bfloat16)
        dout = torch.randn_like(out)
        lse = torch.empty(b, h, t, device="cuda")
        dqkv = torch.empty_like(qkv)
        scratch = torch.empty(K.attention_bwd_scratch_numel(b, t, h), device="cuda")
        flops = 2 * 2 * b * h * t * t * 64 / 2      # causal: half of the score matrix
        for name, fn, mult in (("fwd", lambda: K.attention_fwd(qkv, out, lse, b, t, h), 1.0),
                               ("bwd", lambda: K.attention_bwd(qkv, out, dout, lse, dqkv, scratch, b, t, h), 2.5)):
            for _ in range(3):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(20):
                fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 20
            print(f"  attention {name} b={b}: {ms * 1e3:.1f} us  {flops * mult / ms / 1e9:.0f} TFLOP/s (causal flops)")
    return ok


def group_gemm_perf():
    import torch
    from jax_distributed_tuts_b200.ops import kernels as K
    print("  shape (m,n,k) layout        ours ms   TFLOP/s | torch ms  TFLOP/s")
    for (m, n, k) in [(8192, 8192, 8192), (8192, 3072, 768), (8192, 768, 3072), (8192, 2304, 768),
                      (8192, 768, 768), (16384, 4096, 1024), (8192, 50264, 768)]:
        a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
        bt = torch.randn(n, k, device="cuda", dtype=torch.bfloat16)
        bn_ = torch.randn(k, n, device="cuda", dtype=torch.bfloat16)
        out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * m * n * k
        for name, fn, tfn in (
            ("TN", lambda: K.gemm(a, bt, out), lambda: torch.matmul(a, bt.t(), out=out)),
            ("NN", lambda: K.gemm(a, bn_, out, b_mn=True), lambda: torch.matmul(a, bn_, out=out)),
        ):
            t1 = _time(fn); t2 = _time(tfn)
            print(f"  ({m},{n},{k}) {name}: {t1:8.3f} {fl / t1 / 1e9:8.1f} | {t2:8.3f} {fl / t2 / 1e9:8.1f}", flush=True)
        for bn in (128, 256):
            for pm, nm in ((1, "1-CTA"), (2, "pair ")):
                t1 = _time(lambda: K.gemm(a, bt, out, block_n=bn, pair_mode=pm))
                print(f"     TN block_n={bn} {nm}: {t1:8.3f} ms {fl / t1 / 1e9:8.1f} TFLOP/s", flush=True)
    # wgrad layout
    m, n, k = 768, 3072, 8192
    at = torch.randn(k, m, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(k, n, device="cuda", dtype=torch.bfloat16)
    acc = torch.zeros(m, n, device="cuda")
    t1 = _time(lambda: K.gemm(at, b, acc, a_mn=True, b_mn=True, accumulate=True))
    t2 = _time(lambda: torch.matmul(at.t(), b))
    fl = 2.0 * m * n * k
    print(f"  wgrad ({m},{n},{k}): {t1:8.3f} ms {fl / t1 / 1e9:8.1f} TFLOP/s | torch {t2:8.3f} ms {fl / t2 / 1e9:8.1f}")
    return True


def group_gemm_epi_perf():
    """Cost of each epilogue feature on a GPT-2-shaped GEMM (M=8192, N=3072, K=768)."""
    import torch
    from jax_distributed_tuts_b200.ops import kernels as K
    m, n, k = 8192, 3072, 768
    a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
    bt = torch.randn(n, k, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    out32 = torch.zeros(m, n, device="cuda", dtype=torch.float32)
    pre = torch.randn(m, n, device="cuda", dtype=torch.bfloat16)
    res = torch.randn(m, n, device="cuda", dtype=torch.bfloat16)
    bias = torch.randn(n, device="cuda", dtype=torch.bfloat16)
    fl = 2.0 * m * n * k
    cases = {
        "plain bf16 out": lambda: K.gemm(a, bt, out),
        "plain fp32 out": lambda: K.gemm(a, bt, out32),
        "fp32 accumulate": lambda: K.gemm(a, bt, out32, accumulate=True),
        "+bias": lambda: K.gemm(a, bt, out, bias=bias),
        "+bias+gelu": lambda: K.gemm(a, bt, out, bias=bias, act="gelu"),
        "+bias+gelu+save pre": lambda: K.gemm(a, bt, out, bias=bias, act="gelu", pre=pre),
        "+bias+residual": lambda: K.gemm(a, bt, out, bias=bias, resid=res),
        "bwd aux (no act)": lambda: K.gemm(a, bt, out, bwd=True, aux=pre),
        "bwd aux gelu'": lambda: K.gemm(a, bt, out, bwd=True, aux=pre, act="gelu"),
        "bwd aux silu'": lambda: K.gemm(a, bt, out, bwd=True, aux=pre, act="silu"),
    }
    for name, fn in cases.items():
        t = _time(fn)
        print(f"  {name:24s} {t * 1e3:8.1f} us  {fl / t / 1e9:7.1f} TFLOP/s", flush=True)
    return True


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--group", default=None)
    ap.add_argument("--groups", default=",".join(GROUPS))
    ap.add_argument("--timeout", type=int, default=240)
    a = ap.parse_args()
    if a.group:
        import torch
        assert torch.cuda.is_available(), "needs a GPU"
        t0 = time.time()
        print(f"== group {a.group}", flush=True)
        try:
            ok = globals()["group_" + a.group]()
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001
            import traceback
            traceback.print_exc()
            print(f"  [FAIL] exception: {type(e).__name__}: {e}")
            ok = False
        print(f"== group {a.group}: {'PASS' if ok else 'FAIL'} ({time.time() - t0:.1f}s)", flush=True)
        sys.exit(0 if ok else 1)
    os.makedirs("gpurun_out", exist_ok=True)
    results = {}
    with open("gpurun_out/kernels_check.log", "a") as log:
        for g in a.groups.split(","):
            try:
                r = subprocess.run([sys.executable, __file__, "--group", g], capture_output=True, text=True,
                                   timeout=a.timeout)
                out, rc = r.stdout + r.stderr[-3000:], r.returncode
            except subprocess.TimeoutExpired as e:
                out = (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or "")
                out += "\n  [FAIL] TIMEOUT"
                rc = 124
            results[g] = rc
            print(out, flush=True)
            log.write(out + "\n")
        summary = "SUMMARY " + " ".join(f"{g}={'ok' if rc == 0 else 'FAIL(%d)' % rc}" for g, rc in results.items())
        print(summary)
        log.write(summary + "\n")
    sys.exit(0 if all(rc == 0 for rc in results.values()) else 1)


if __name__ == "__main__":
    main()

import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from jax_distributed_tuts_b200.ops import kernels as K
torch.manual_seed(0)
for (m, n, k, kw) in [(64, 512, 16, dict()), (64, 512, 16, dict(bwd=True, act="silu")), (300, 768, 256, dict(bwd=True, act="gelu")),
                      (8192, 3072, 768, dict(bwd=True, act="gelu"))]:
    a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(n, k, device="cuda", dtype=torch.bfloat16) * 0.1
    aux = torch.randn(m, n, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    db = torch.zeros(n, device="cuda")
    key = torch.tensor([5], device="cuda", dtype=torch.int64)
    extra = dict(aux=aux) if kw.get("bwd") else {}
    for drop in (0.0, 0.1):
        db.zero_()
        K.gemm(a, b, out, dbias=db, drop_key=key, drop_rate=drop, drop_salts=[3, 4], drop_rows_per_salt=max(1, m // 2), **kw, **extra)
        ref = out.float().sum(0)
        err = (db - ref).abs().max().item() / ref.abs().max().item()
        print(f"m={m} n={n} k={k} {kw} drop={drop}: dbias rel err {err:.2e}  (|ref|max {ref.abs().max().item():.2f})", flush=True)

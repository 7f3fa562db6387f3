"""Tiny driver for ncu: a few launches of the tcgen05 GEMM. usage: prof_gemm.py {big|gpt} [pair_mode]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from jax_distributed_tuts_b200.ops import kernels as K

shape = sys.argv[1] if len(sys.argv) > 1 else "big"
pm = int(sys.argv[2]) if len(sys.argv) > 2 else 0
m, n, k = (8192, 8192, 8192) if shape == "big" else (8192, 3072, 768)
a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
b = torch.randn(n, k, device="cuda", dtype=torch.bfloat16)
out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
for _ in range(5):
    K.gemm(a, b, out, pair_mode=pm, block_n=256)
torch.cuda.synchronize()
print("ok", out.float().abs().mean().item())

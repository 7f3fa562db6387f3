"""Probe: throughput of the four operand-major combinations on a large GEMM, and wgrad-shaped splits."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from jax_distributed_tuts_b200.ops import kernels as K
from tools.gpu_check_kernels import _time

m = n = k = 8192
a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16); at = a.t().contiguous()
b = torch.randn(k, n, device="cuda", dtype=torch.bfloat16); bt = b.t().contiguous()
out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
fl = 2.0 * m * n * k
for name, A, B, amn, bmn in (("A K-major, B K-major", a, bt, False, False), ("A K-major, B MN-major", a, b, False, True),
                             ("A MN-major, B K-major", at, bt, True, False), ("A MN-major, B MN-major", at, b, True, True)):
    for pm in (1, 2):
        t = _time(lambda: K.gemm(A, B, out, a_mn=amn, b_mn=bmn, pair_mode=pm, block_n=256), iters=8)
        print(f"{name:24s} pair_mode={pm}: {t:7.3f} ms {fl / t / 1e9:7.1f} TFLOP/s", flush=True)
# wgrad shape with different split counts
m, n, k = 768, 3072, 8192
at = torch.randn(k, m, device="cuda", dtype=torch.bfloat16); b = torch.randn(k, n, device="cuda", dtype=torch.bfloat16)
acc = torch.zeros(m, n, device="cuda")
fl = 2.0 * m * n * k
for pm in (1, 2):
    for bn in (128, 256):
        for sk in (1, 2, 4, 8):
            t = _time(lambda: K.gemm(at, b, acc, a_mn=True, b_mn=True, pair_mode=pm, block_n=bn, split_k=sk, accumulate=True), iters=8)
            print(f"wgrad pair_mode={pm} bn={bn} split_k={sk}: {t * 1e3:7.1f} us {fl / t / 1e9:7.1f} TFLOP/s", flush=True)

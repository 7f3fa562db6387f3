"""One launch of every kernel family on GPT-2-small shapes (M = 8192 tokens), for a single ncu capture."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from jax_distributed_tuts_b200.ops import kernels as K

torch.manual_seed(0)
M, C, H, T, B = 8192, 768, 12, 1024, 8
bf = torch.bfloat16
x = torch.randn(M, C, device="cuda", dtype=bf)
w_fc = torch.randn(C, 4 * C, device="cuda", dtype=bf) * 0.02
b_fc = torch.zeros(4 * C, device="cuda", dtype=bf)
hpre = torch.empty(M, 4 * C, device="cuda", dtype=bf); hact = torch.empty_like(hpre)
dx = torch.randn(M, C, device="cuda", dtype=bf)
w_pr = torch.randn(4 * C, C, device="cuda", dtype=bf) * 0.02
dh = torch.empty(M, 4 * C, device="cuda", dtype=bf)
gw = torch.zeros(C, 4 * C, device="cuda")
gb = torch.zeros(4 * C, device="cuda")
for rep in range(2):   # first pass warms up (module load, attributes); ncu is told to skip it
    K.gemm(x, w_fc, hact, b_mn=True, bias=b_fc, act="gelu", pre=hpre)                  # fwd fc: pair kernel, EPI_BIAS_GELU_PRE
    K.gemm(dx, w_pr, dh, bwd=True, aux=hpre, act="gelu")                               # dgrad + GELU'
    K.gemm(x, dh, gw, a_mn=True, b_mn=True)                                            # wgrad, split-K
    K.colsum(dh, gb)
    g = torch.ones(C, device="cuda", dtype=bf); bt = torch.zeros(C, device="cuda", dtype=bf)
    y = torch.empty_like(x); mean = torch.empty(M, device="cuda"); rstd = torch.empty(M, device="cuda")
    K.layernorm_fwd(x, g, bt, y, mean, rstd)
    dg = torch.zeros(C, device="cuda"); db = torch.zeros(C, device="cuda"); dxo = torch.empty_like(x)
    K.layernorm_bwd(dx, x, g, mean, rstd, dxo, dg, db, dresid=dx)
    qkv = torch.randn(B, T, 3, H, 64, device="cuda", dtype=bf)
    out = torch.empty(B, T, H, 64, device="cuda", dtype=bf); lse = torch.empty(B, H, T, device="cuda")
    K.attention_fwd(qkv, out, lse, B, T, H)
    dqkv = torch.empty_like(qkv); scratch = torch.empty(K.attention_bwd_scratch_numel(B, T, H), device="cuda")
    K.attention_bwd(qkv, out, torch.randn_like(out), lse, dqkv, scratch, B, T, H)
    logits = torch.randn(M, 50304, device="cuda", dtype=bf)
    K.softmax_xent_bf16_(logits[:, :50257], torch.randint(0, 50257, (M,), device="cuda"), grad_scale=1.0 / M,
                         metrics=torch.zeros(2, device="cuda"))
    n = 7_087_872
    p = torch.randn(n, device="cuda"); m_ = torch.zeros(n, device="cuda"); v_ = torch.zeros(n, device="cuda")
    K.adamw_(p, m_, v_, torch.randn(n, device="cuda"), torch.empty(n, device="cuda", dtype=bf),
             torch.zeros(1, device="cuda", dtype=torch.int32), lr=1e-3)
    torch.cuda.synchronize()
print("ok")

"""Probe: does a TMA load whose inner coordinate is not 16-byte aligned fault? (K-sharded GEMM)"""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

def case(m, n, k, p):
    import torch
    from jax_distributed_tuts_b200.ops import kernels as K
    torch.manual_seed(0)
    x = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
    w = (torch.randn(k, n, device="cuda") / k ** 0.5).bfloat16()
    ref = x.float() @ w.float()
    torch.cuda.synchronize()
    shards = [s.contiguous() for s in w.chunk(p, dim=0)]
    out = torch.full((m, n), float("nan"), device="cuda", dtype=torch.float32)
    K.gemm(x, shards, out, b_mn=True, shard_mode=1, shard_len=k // p, rot=1 % p, b_shape=(k, n), ldb=n)
    torch.cuda.synchronize()
    print(f"m={m} n={n} k={k} p={p} shard_len={k//p}: rel_err={(out-ref).abs().max().item()/ref.abs().max().item():.3e}", flush=True)

if len(sys.argv) > 1:
    case(*map(int, sys.argv[1:]))
else:
    for c in [(256, 256, 1024, 4), (128, 512, 832, 8), (128, 512, 800, 8), (64, 512, 784, 8)]:
        r = subprocess.run([sys.executable, __file__, *map(str, c)], capture_output=True, text=True, timeout=200)
        print(r.stdout.strip() or ("FAIL " + str(c) + " " + r.stderr.strip().splitlines()[-1]))

"""One small launch of every kernel family on odd shapes (partial tiles, short last row group,
item switches in the persistent attention kernels): a quick "does everything still launch and
finish" check, e.g. as the target of a memory checker where one is available.
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from jax_distributed_tuts_b200.ops import kernels as K

torch.manual_seed(0)
bf = torch.bfloat16
dev = "cuda"

# Guard bands: with --guard every tensor this script creates sits in the middle of a larger allocation
# whose 4 KiB on either side hold a byte pattern; after all kernels have run, every band must still hold
# it. A kernel that stores outside the tensor it was given (partial tiles, odd pitches, tails) is caught
# without a memory checker (compute-sanitizer is closed on this GPU pool).
GUARD = "--guard" in sys.argv
_BANDS = []
_PAD = 4096


def _guarded(t: torch.Tensor) -> torch.Tensor:
    if not GUARD or not t.is_cuda:
        return t
    nbytes = t.numel() * t.element_size()
    nb16 = (nbytes + 15) // 16 * 16
    raw = torch.full((_PAD + nb16 + _PAD,), 0xA5, dtype=torch.uint8, device=t.device)
    mid = raw[_PAD:_PAD + nbytes].view(t.dtype).view(t.shape)
    mid.copy_(t)
    _BANDS.append((raw, nbytes))
    return mid


for _name in ("randn", "empty", "zeros", "ones", "randint", "randn_like", "empty_like", "zeros_like"):
    def _wrap(fn):
        def inner(*a, **k):
            return _guarded(fn(*a, **k))
        return inner
    setattr(torch, _name, _wrap(getattr(torch, _name)))


def check_guards() -> int:
    bad = 0
    for raw, nbytes in _BANDS:
        lo, hi = raw[:_PAD], raw[_PAD + nbytes:]
        if not bool((lo == 0xA5).all()) or not bool((hi == 0xA5).all()):
            bad += 1
    return bad
# GEMMs: plain, bias+gelu+pre, bwd gelu' with dbias, wgrad fp32 (split-K), fused cross-entropy head
m, n, k = 256, 384, 320
x = torch.randn(m, k, device=dev, dtype=bf); w = torch.randn(k, n, device=dev, dtype=bf) * 0.05
bias = torch.randn(n, device=dev, dtype=bf); out = torch.empty(m, n, device=dev, dtype=bf); pre = torch.empty_like(out)
K.gemm(x, w, out, b_mn=True)
K.gemm(x, w, out, b_mn=True, bias=bias, act="gelu", pre=pre)
dy = torch.randn(m, n, device=dev, dtype=bf); dx = torch.empty(m, k, device=dev, dtype=bf); db = torch.zeros(k + 8, device=dev)
aux = torch.randn(m, k, device=dev, dtype=bf)
K.gemm(dy, w, dx, bwd=True, aux=aux, act="gelu", dbias=db[:k])
gw = torch.zeros(k, n, device=dev)
K.gemm(x, dy, gw, a_mn=True, b_mn=True)
wh = (torch.randn(k, 16, device=dev) * 0.05).to(bf)[:, :10]; dl = torch.zeros(m, 16, device=dev, dtype=bf)[:, :10]
K.gemm(x, wh, dl, b_mn=True, bias=bias[:10], xent_labels=torch.randint(0, 10, (m,), device=dev),
       xent_metrics=torch.zeros(2, device=dev), xent_scale=1.0 / m, dbias=torch.zeros(16, device=dev))
# element-wise / transformer kernels
rows, cols = 70, 768
h = torch.randn(rows, cols, device=dev, dtype=bf); g = torch.ones(cols, device=dev, dtype=bf); b = torch.zeros(cols, device=dev, dtype=bf)
y = torch.empty_like(h); mean = torch.empty(rows, device=dev); rstd = torch.empty(rows, device=dev)
K.layernorm_fwd(h, g, b, y, mean, rstd)
K.layernorm_bwd(torch.randn_like(h), h, g, mean, rstd, torch.empty_like(h), torch.zeros(cols, device=dev), torch.zeros(cols, device=dev),
                dresid=torch.randn_like(h))
logits = torch.randn(9, 50264, device=dev, dtype=bf)[:, :50257]
K.softmax_xent_bf16_(logits, torch.randint(0, 50257, (9,), device=dev), grad_scale=0.1, metrics=torch.zeros(2, device=dev))
small = torch.randn(33, 10, device=dev)
K.softmax_xent(small, torch.randint(0, 10, (33,), device=dev), grad_scale=1.0, dlogits=torch.empty(33, 16, device=dev, dtype=bf)[:, :10],
               metrics=torch.zeros(2, device=dev))
nparam = 4096 * 3 + 8
p = torch.randn(nparam, device=dev); K.adamw_(p, torch.zeros_like(p), torch.zeros_like(p), torch.randn_like(p),
                                               torch.empty(nparam, device=dev, dtype=bf), torch.zeros(1, device=dev, dtype=torch.int32), lr=1e-3)
# attention forward + backward (two sequences, three heads, two q tiles: every pipeline path incl. item switches)
B, T, H = 2, 256, 3
qkv = torch.randn(B, T, 3, H, 64, device=dev, dtype=bf); o = torch.empty(B, T, H, 64, device=dev, dtype=bf); lse = torch.empty(B, H, T, device=dev)
K.attention_fwd(qkv, o, lse, B, T, H)
K.attention_bwd(qkv, o, torch.randn_like(o), lse, torch.empty_like(qkv), torch.empty(K.attention_bwd_scratch_numel(B, T, H), device=dev), B, T, H)
torch.cuda.synchronize()
print("all kernels launched and finished")
if GUARD:
    bad = check_guards()
    print(f"guard bands: {len(_BANDS)} tensors, {bad} overrun(s)")
    sys.exit(1 if bad else 0)

"""Ablation timing of the attention kernels (development tool).

`--build` (no GPU needed) compiles csrc/attention.cu with -DATT_EXP=<mask> for each mask in
VARIANTS and links it with the other objects into jax_distributed_tuts_b200/_lib/exp/;
`--run` (on a B200) times forward and backward of every variant in a fresh subprocess.
"""
import os
import subprocess
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))
EXP_DIR = REPO / "jax_distributed_tuts_b200" / "_lib" / "exp"
VARIANTS = [int(v) for v in os.environ.get("ATT_EXP_VARIANTS", "0").split(",")]


def build():
    from jax_distributed_tuts_b200.ops import _build
    _build.build()
    EXP_DIR.mkdir(parents=True, exist_ok=True)
    others = [str(_build.OBJ_DIR / (s.stem + ".o")) for s in _build._sources() if s.stem != "attention"]
    for v in VARIANTS:
        obj = _build.OBJ_DIR / f"attention_exp{v}.o"
        subprocess.run([_build._nvcc(), *_build.NVCC_FLAGS, f"-DATT_EXP={v}", *(["-DATT_TRACE"] if v & 16 else []), "-c", str(_build.CSRC / "attention.cu"),
                        "-o", str(obj)], check=True)
        subprocess.run([_build._nvcc(), "-shared", "-o", str(EXP_DIR / f"libexp{v}.so"), str(obj), *others], check=True)
        print("built variant", v)


def child():
    import torch
    from jax_distributed_tuts_b200.ops import kernels as K
    from torch.profiler import ProfilerActivity, profile
    b, t, h = 16, 1024, 12
    qkv = torch.randn(b, t, 3, h, 64, device="cuda").bfloat16()
    out = torch.empty(b, t, h, 64, device="cuda", dtype=torch.bfloat16)
    dout = torch.randn_like(out)
    lse = torch.empty(b, h, t, device="cuda")
    dqkv = torch.empty_like(qkv)
    scratch = torch.empty(K.attention_bwd_scratch_numel(b, t, h), device="cuda")
    for _ in range(3):
        K.attention_fwd(qkv, out, lse, b, t, h)
        K.attention_bwd(qkv, out, dout, lse, dqkv, scratch, b, t, h)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(10):
            K.attention_fwd(qkv, out, lse, b, t, h)
            K.attention_bwd(qkv, out, dout, lse, dqkv, scratch, b, t, h)
        torch.cuda.synchronize()
    parts = []
    for ev in prof.key_averages():
        if "attention" in ev.key:
            name = ev.key.split("(")[0].replace("b200::attention_", "")
            parts.append(f"{name}={ev.device_time_total / max(ev.count, 1):.1f}us")
    print("  ".join(parts), flush=True)


def trace():
    """Cycle stamps of one late backward CTA (library built with -DATT_TRACE)."""
    import torch
    from jax_distributed_tuts_b200.ops import kernels as K
    b, t, h = 16, 1024, 12
    qkv = torch.randn(b, t, 3, h, 64, device="cuda").bfloat16()
    out = torch.empty(b, t, h, 64, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(b, h, t, device="cuda")
    K.attention_fwd(qkv, out, lse, b, t, h)
    dout = torch.randn_like(out)
    dqkv = torch.empty_like(qkv)
    n = K.attention_bwd_scratch_numel(b, t, h)
    scratch = torch.zeros(n + 4096, device="cuda")
    for _ in range(2):
        K.attention_bwd(qkv, out, dout, lse, dqkv, scratch[:n], b, t, h)
    torch.cuda.synchronize()
    tr = scratch[n:].cpu().tolist()
    print("bwd math thread per item: [item start, first sdp_full, 2nd step top, loop done, last dq_full, flushed, dK/dV stored, niter]")
    for j in range(12):
        print("  n=%2d " % j + " ".join("%7.0f" % tr[j * 8 + k] for k in range(8)))
    print("bwd MMA thread per item: [item start, kv_full seen, first S/dP issued]")
    for j in range(12):
        print("  n=%2d " % j + " ".join("%7.0f" % tr[512 + j * 4 + k] for k in range(3)))


def run():
    for v in VARIANTS:
        env = dict(os.environ, B200DIST_LIB=str(EXP_DIR / f"libexp{v}.so"))
        r = subprocess.run([sys.executable, __file__, "--child"], env=env, capture_output=True, text=True, timeout=300)
        tail = (r.stdout.strip().splitlines() or ["?"])[-1]
        print(f"variant {v:2d}: {tail}"
              + ("" if r.returncode == 0 else f"  [rc={r.returncode}] {r.stderr[-300:]}"), flush=True)


if __name__ == "__main__":
    if "--build" in sys.argv:
        build()
    elif "--child" in sys.argv:
        child()
    elif "--trace" in sys.argv:
        trace()
    else:
        run()

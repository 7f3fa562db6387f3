"""A/B timing of the attention kernels across library builds (development tool, run on a B200).

    python tools/exp_attention.py [path/to/libA.so path/to/libB.so ...]

Each library (default: the in-tree build) is loaded in a fresh subprocess through B200DIST_LIB and
forward + backward are timed per kernel with CUPTI (torch.profiler) at the GPT-2 small bench shape,
16 x 1024 tokens, 12 heads. Alternative builds are made by compiling a modified csrc/attention.cu
and linking it with the other objects of build/obj (see ops/_build.py); this is how the ablations
quoted in DESIGN.md ("Measured findings") were run.
"""
import os
import subprocess
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))


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


def run(libs):
    for lib in libs:
        env = dict(os.environ, B200DIST_NO_PDL="1")      # kernel times without dependent-launch overlap
        if lib:
            env["B200DIST_LIB"] = str(Path(lib).resolve())
        r = subprocess.run([sys.executable, __file__, "--child"], env=env, capture_output=True, text=True, timeout=300)
        tail = (r.stdout.strip().splitlines() or ["?"])[-1]
        print(f"{lib or 'in-tree build'}: {tail}" + ("" if r.returncode == 0 else f"  [rc={r.returncode}] {r.stderr[-300:]}"),
              flush=True)


if __name__ == "__main__":
    if "--child" in sys.argv:
        child()
    else:
        run([a for a in sys.argv[1:] if not a.startswith("--")] or [None])

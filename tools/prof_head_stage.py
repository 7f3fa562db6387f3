"""Kernel table of a head-only pipeline stage (final norm + tied LM head + loss, forward and backward) for one
microbatch of GPT-2 small: where the last stage of gpt2s_pp4 spends its time. One GPU."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile
from jax_distributed_tuts_b200 import rng
from jax_distributed_tuts_b200.engine import GPT2Engine
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.gpt2 import GPT2, gpt2_config

cfg = gpt2_config("small")
model = GPT2(cfg)
mesh = Mesh((1,), ("data",))
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 2
with mesh:
    params = model.init({"params": rng.PRNGKey(1)}, torch.zeros(1, 1024, dtype=torch.int64))["params"]
    eng = GPT2Engine(cfg, mesh, mode="dp", batch_per_rank=4 * mb, seq_len=1024, lr=3e-4, num_minibatches=4, init_params=params,
                     layer_ids=[], has_embed=False, has_head=True, slots=4, use_graph=False, overlap=False)
    eng.targets.random_(0, cfg.vocab_size)
    for s in eng.slot_x0:
        s.normal_()

    def one(k):
        eng._forward(None, slot=k)
        eng._head_backward(eng.targets[k * mb:(k + 1) * mb], eng.slot_dx_top[k], accumulate=k > 0)

    for k in range(4):
        one(k)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for k in range(4):
        one(k)
    b.record()
    torch.cuda.synchronize()
    print(f"head stage, {mb} sequences per microbatch: {a.elapsed_time(b) / 4:.3f} ms per microbatch (eager)")
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for k in range(4):
            one(k)
        torch.cuda.synchronize()
    for ev in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:12]:
        print(f"  {ev.device_time_total / 4e3:8.3f} ms  x{ev.count // 4:3d}  {ev.key[:120]}")

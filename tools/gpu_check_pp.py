"""Pipeline engine check (torchrun, N = 2, 4 or 8 GPUs): the GPipe engine over S stages (and D data
replicas when N > S) must reproduce the single-GPU fused GPT2Engine run with the same microbatches:
per-step losses and, per stage, the parameters it owns after three steps.

    torchrun --nproc-per-node 4 tools/gpu_check_pp.py [--stages 4]
"""
from __future__ import annotations

import argparse
import os
import sys
import traceback

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from jax_distributed_tuts_b200 import rng
from jax_distributed_tuts_b200.engine import GPT2Engine
from jax_distributed_tuts_b200.engine.pp import PipelineEngine
from jax_distributed_tuts_b200.launch import init_from_env
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.gpt2 import GPT2, gpt2_config
from jax_distributed_tuts_b200.state import Batch
from jax_distributed_tuts_b200.tree import tree_leaves_with_path


def log(*a):
    if dist.get_rank() == 0:
        print(*a, flush=True)


def run_case(rank, world, device, S, use_graph, schedule="1f1b") -> bool:
    D = world // S
    cfg = gpt2_config("tiny", n_layer=4, d_model=256, n_head=4, vocab_size=1024, n_ctx=256)
    model = GPT2(cfg)
    B, T, Mb, steps, lr = 8, 128, 4, 3, 1e-3
    g = torch.Generator().manual_seed(0)
    tok = torch.randint(0, cfg.vocab_size, (B * D, T + 1), generator=g)
    params = model.init({"params": rng.PRNGKey(3)}, tok[:1, :T])["params"]
    # ---- oracle: the full model on this GPU alone, same microbatches, the whole global batch ----
    solo = Mesh((1,), ("data",), rank=0, world_size=1, device=device)
    with solo:
        ref = GPT2Engine(cfg, solo, mode="dp", batch_per_rank=B * D, seq_len=T, lr=lr, num_minibatches=Mb * D,
                         init_params=params, use_graph=False)
        st = ref.make_state(model, rng.PRNGKey(5))
        gb = Batch(tok[:, :T].contiguous().pin_memory(), tok[:, 1:].contiguous().pin_memory())
        ref_losses = []
        for _ in range(steps):
            st, m = ref.train_step(st, None, gb)
            ref_losses.append(float(m["loss"][0]) / m["loss"][1])
        torch.cuda.synchronize()
        ref_params = dict(tree_leaves_with_path(ref.params_tree()))
    del ref
    mesh = Mesh((D, S), ("data", "pipe"))
    d = mesh.axis_index("data")
    lb = Batch(tok[d * B:(d + 1) * B, :T].contiguous().pin_memory(), tok[d * B:(d + 1) * B, 1:].contiguous().pin_memory())
    with mesh:
        eng = PipelineEngine(cfg, mesh, batch_per_replica=B, seq_len=T, num_microbatches=Mb, lr=lr, init_params=params,
                             use_graph=use_graph, schedule=schedule)
        losses = []
        for _ in range(steps):
            ls, n = eng.train_step(lb)
            losses.append(float(ls) / n)
        torch.cuda.synchronize()
        got = dict(tree_leaves_with_path(eng.params_tree()))
        # the merged checkpoint holds the whole model and reloads into the stages bit for bit
        sd = eng.state_dict()
        want_keys = sorted([f"h_{i}" for i in range(cfg.n_layer)] + ["wte", "wpe", "ln_f"])
        ckpt_ok = sorted(sd["params"].keys()) == want_keys and sorted(sd["opt_state"]["0"]["mu"].keys()) == want_keys
        before = {k: v.clone() for k, v in got.items()}
        eng.load_state_dict(sd)
        torch.cuda.synchronize()
        again = dict(tree_leaves_with_path(eng.params_tree()))
        ckpt_ok &= all(torch.equal(before[k], again[k]) for k in before)
    ok = ckpt_ok and all(abs(a - b) < 2e-3 for a, b in zip(ref_losses, losses))
    worst = ("", 0.0)
    for k, v in got.items():
        r = ref_params[k]
        dmean = (v.float() - r.float()).abs().mean().item()
        if dmean > worst[1]:
            worst = (k, dmean)
        # Adam moves every element by about lr per step: a wrong or missing gradient shows up at that scale
        if dmean > 0.08 * lr * steps:
            ok = False
    t = torch.tensor([int(ok)], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    log(f"[{'ok' if t.item() else 'FAIL'}] pipeline S={S} D={D} graphs={use_graph} {schedule}: losses {[f'{x:.4f}' for x in losses]} "
        f"(single GPU {[f'{x:.4f}' for x in ref_losses]}) split {[len(r) for r in eng.split]} "
        f"worst leaf on stage 0: {worst[0]} mean|d|={worst[1]:.2e}")
    del eng
    torch.cuda.empty_cache()
    return bool(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stages", type=int, default=0)
    a = ap.parse_args()
    rank, world, device = init_from_env()
    ok = True
    cases = [(a.stages, True, "1f1b")] if a.stages else \
        [(world, False, "gpipe"), (world, True, "gpipe"), (world, False, "1f1b"), (world, True, "1f1b")] + \
        ([(world // 2, True, "1f1b")] if world >= 4 else [])
    for S, graph, sched in cases:
        try:
            ok &= run_case(rank, world, device, S, graph, sched)
        except Exception as e:  # noqa: BLE001
            traceback.print_exc()
            log(f"[FAIL] pipeline S={S}: {type(e).__name__}: {e}")
            ok = False
            break
    log("PIPELINE CHECK", "PASS" if ok else "FAIL")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()

"""Per-phase timeline of the fused reduce + AdamW + publish kernel inside a real multi-GPU step.

    torchrun --nproc-per-node N tools/comm_phase_timeline.py [--config mlp_dp|mlp_fsdp] [--out FILE]

Thread 0 of every block stores %globaltimer (ns) at five points (b200_comm_set_stamps): kernel start,
after the "gradients ready" flag exchange, after the reduce/AdamW/publish loop, after the system
fence + "published" increments, at the end. The step runs as the usual CUDA graph with a 256 MiB L2
flush in front of every step; the table is the median over steps of (per-rank max over blocks) and
then the max over ranks, in microseconds.
"""
from __future__ import annotations

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from jax_distributed_tuts_b200 import rng
from jax_distributed_tuts_b200.config import dp_config
from jax_distributed_tuts_b200.engine import MLPEngine
from jax_distributed_tuts_b200.launch import init_from_env
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.mlp import Classifier
from jax_distributed_tuts_b200.ops import _lib
from jax_distributed_tuts_b200.parallel import dp
from jax_distributed_tuts_b200.state import Batch


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="mlp_dp", choices=["mlp_dp", "mlp_fsdp"])
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, world, device = init_from_env()
    cfg = dp_config()
    mesh = Mesh((world,), ("data",))
    model = Classifier(cfg.model)
    key = rng.PRNGKey(cfg.seed)
    model_rng, data_rng, label_rng = rng.split(key, 3)
    x = torch.randn(16, 784, generator=rng.uniform_generator(rng.fold_in(data_rng, rank))).pin_memory()
    y = torch.randint(0, 10, (16,), generator=rng.uniform_generator(rng.fold_in(label_rng, rank))).pin_memory()
    mode = "fsdp" if a.config == "mlp_fsdp" else "dp"
    stamps = torch.zeros(_lib.COMM_MAX_BLOCKS * 8, dtype=torch.int64, device=device)
    with mesh:
        init = dp.init_dp(model_rng, x[:4], model)
        eng = MLPEngine(cfg.model, mesh, mode=mode, batch_per_rank=16, lr=1e-3, num_minibatches=4,
                        init_params=init.params, min_weight_size=16 if mode == "fsdp" else 2 ** 18)
        st = eng.make_state(model, init.rng)
        batch = Batch(x, y)
        _lib.lib().b200_comm_set_stamps(stamps.data_ptr())       # read at capture time: baked into the graph
        for _ in range(4):
            st, _m = eng.train_step(st, None, batch)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
        rows, step_ms = [], []
        for _ in range(a.steps):
            flush.zero_()
            eng._load_batch(batch)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); eng._run_step(); e.record()
            torch.cuda.synchronize()
            t = stamps.view(-1, 8)[:, :5].cpu()
            t = t[t[:, 0] > 0]
            t0 = t[:, 0].min()
            # phase ends, max over blocks, relative to the first block's start
            rows.append([(t[:, k].max() - t0).item() / 1e3 for k in range(5)] + [(t[:, 0].max() - t0).item() / 1e3])
            step_ms.append(s.elapsed_time(e))
        _lib.lib().b200_comm_set_stamps(None)
        r = torch.tensor(rows, dtype=torch.float64).median(0).values
        ms = torch.tensor(step_ms, dtype=torch.float64).median().to(device)
        rr = r.to(device)
        if world > 1:
            dist.all_reduce(rr, op=dist.ReduceOp.MAX)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if rank == 0:
            last_start, ready, upd, pub, end = rr[5].item(), rr[1].item(), rr[2].item(), rr[3].item(), rr[4].item()
            lines = [
                f"| {a.config} | {world} | {eng.arena.backend} | {ms.item() * 1e3:.1f} | {last_start:.1f} | {ready:.1f} | "
                f"{upd - ready:.1f} | {pub - upd:.1f} | {end:.1f} |"]
            hdr = ("| config | GPUs | comm | step us (median, max over ranks) | last block starts | flag exchange done | "
                   "reduce+AdamW+publish | fence.sys + published | kernel total |\n|---|---|---|---|---|---|---|---|---|")
            print(hdr)
            print("\n".join(lines), flush=True)
            if a.out:
                new = not os.path.exists(a.out)
                with open(a.out, "a") as f:
                    if new:
                        f.write(hdr + "\n")
                    f.write("\n".join(lines) + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""Multi-GPU GPT-2 check (torchrun, one rank per GPU): fused DP and FSDP engines (tiny GPT-2 so the
generic NCCL path is a cheap oracle) must agree with generic data parallelism, replicas must be
identical; then GPT-2 small step times for DP and FSDP."""
from __future__ import annotations

import os
import sys
import traceback

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from jax_distributed_tuts_b200 import rng
from jax_distributed_tuts_b200.engine import GPT2Engine
from jax_distributed_tuts_b200.launch import init_from_env
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.gpt2 import GPT2, gpt2_config, lm_loss_fn
from jax_distributed_tuts_b200.optim import adamw
from jax_distributed_tuts_b200.parallel import dp
from jax_distributed_tuts_b200.state import Batch, TrainState
from jax_distributed_tuts_b200.tree import tree_leaves_with_path, tree_map


def log(*a):
    if dist.get_rank() == 0:
        print(*a, flush=True)


def main():
    rank, world, device = init_from_env()
    ok_all = True
    cfg = gpt2_config("tiny", n_layer=2, d_model=256, n_head=4, vocab_size=1024, n_ctx=256)
    model = GPT2(cfg)
    per, T, steps, lr = 4, 128, 3, 1e-3
    g = torch.Generator().manual_seed(0)
    tok = torch.randint(0, cfg.vocab_size, (per * world, T + 1), generator=g)
    gb = Batch(tok[:, :T].contiguous(), tok[:, 1:].contiguous())
    mesh = Mesh((world,), ("data",))
    with mesh:
        params = model.init({"params": rng.PRNGKey(3)}, gb.inputs[:1])["params"]
        local = dp.shard_batch(gb)
        pg = tree_map(lambda t: t.to(device), params)
        st = TrainState.create(apply_fn=model.apply, params=pg, tx=adamw(lr), rng=rng.PRNGKey(5).to(device))
        step = dp.make_train_step_dp(mesh, num_minibatches=1, loss_fn=lm_loss_fn)
        lb = local.to(device)
        gen_losses = []
        for _ in range(steps):
            st, m = step(st, None, lb)
            gen_losses.append(float(m["loss"][0]) / m["loss"][1])
        log("generic NCCL DP losses", [f"{x:.4f}" for x in gen_losses])
        ref = dict(tree_leaves_with_path(st.params))
        for name, mode, mws, fw in (("dp", "dp", 2 ** 18, "fused"), ("fsdp/fused", "fsdp", 2 ** 12, "fused"),
                                    ("fsdp/gather", "fsdp", 2 ** 12, "gather")):
            try:
                eng = GPT2Engine(cfg, mesh, mode=mode, batch_per_rank=per, seq_len=T, lr=lr, num_minibatches=1,
                                 init_params=params, use_graph=True, min_weight_size=mws, comm_blocks=32,
                                 fsdp_weights=fw)
                s2 = eng.make_state(model, rng.PRNGKey(5))
                losses = []
                for _ in range(steps):
                    s2, m2 = eng.train_step(s2, None, local)
                    losses.append(float(m2["loss"][0]) / m2["loss"][1])
                torch.cuda.synchronize()
                good = all(abs(a - b) < 3e-2 for a, b in zip(gen_losses, losses))
                got = dict(tree_leaves_with_path(eng.params_tree()))
                worst = ("", 0.0)
                for k, a in ref.items():
                    b = got[k]
                    b = b[: a.shape[0]] if b.shape != a.shape else b
                    d = (a.float() - b.float()).abs()
                    tol = (0.6 if k.endswith("bias") else 0.08) * lr * steps
                    if d.mean().item() > tol or d.max().item() > 2.2 * lr * steps:
                        good = False
                    if d.mean().item() > worst[1]:
                        worst = (k, d.mean().item())
                # Adam second moments, reassembled from the owner-update slices (DP) or the row / column
                # shards (FSDP) into the global per-leaf layout, against the generic path's state
                # (nu = EMA of g^2 is insensitive to the sign noise that makes mu of tiny gradients jitter)
                og = eng.opt_state_global()
                nu_ref = dict(tree_leaves_with_path(st.opt_state[0].nu))
                nu_got = dict(tree_leaves_with_path(og["nu"]))
                worst_nu = ("", 0.0)
                for k, a in nu_ref.items():
                    b = nu_got[k]
                    b = b[: a.shape[0]] if b.shape != a.shape else b
                    if a.float().norm().item() < 1e-9 and b.float().norm().item() < 1e-9:
                        continue          # exactly-zero gradients (attention key bias): only rounding noise
                    err = ((a.float() - b.float()).norm() / a.float().norm().clamp_min(1e-20)).item()
                    if err > worst_nu[1]:
                        worst_nu = (k, err)
                if not worst_nu[1] < 0.15 or int(og["count"].item()) != steps:
                    good = False
                    log(f"  [{name}] optimizer state mismatch: nu {worst_nu[0]} rel err {worst_nu[1]:.3e}, count {int(og['count'].item())}")
                flat = torch.cat([t.reshape(-1).float() for t in got.values()])
                ref0 = flat.clone()
                dist.broadcast(ref0, 0)
                same = torch.tensor([int(torch.equal(ref0, flat))], device=device)
                dist.all_reduce(same, op=dist.ReduceOp.MIN)
                sharded = sum(1 for v in getattr(eng, "shard_dim", {}).values() if v is not None)
                log(f"[{'ok' if good and same.item() else 'FAIL'}] gpt2-tiny {name}: losses {[f'{x:.4f}' for x in losses]} "
                    f"replicas_identical={bool(same.item())} worst leaf {worst[0]} mean|d|={worst[1]:.2e} "
                    f"backend={eng.arena.backend} sharded_leaves={sharded}")
                ok_all &= good and bool(same.item())
                del eng
            except Exception as e:  # noqa: BLE001
                traceback.print_exc()
                log(f"[FAIL] gpt2-tiny {name}: {type(e).__name__}: {e}")
                ok_all = False
        # ---- 2-D mesh (data x fsdp): shards over `fsdp`, gradient mean over both axes ----
        if world >= 4:
            try:
                mesh2 = Mesh((2, world // 2), ("data", "fsdp"))
                with mesh2:
                    eng = GPT2Engine(cfg, mesh2, mode="fsdp", shard_axis="fsdp", batch_per_rank=per, seq_len=T, lr=lr,
                                     num_minibatches=1, init_params=params, use_graph=True, min_weight_size=2 ** 12,
                                     comm_blocks=32, fsdp_weights="fused")
                    s2 = eng.make_state(model, rng.PRNGKey(5))
                    losses = []
                    for _ in range(steps):
                        s2, m2 = eng.train_step(s2, None, local)
                        losses.append(float(m2["loss"][0]) / m2["loss"][1])
                    torch.cuda.synchronize()
                    good = all(abs(a - b) < 3e-2 for a, b in zip(gen_losses, losses))
                    got = dict(tree_leaves_with_path(eng.params_tree()))
                    for k, a in ref.items():
                        b = got[k]
                        b = b[: a.shape[0]] if b.shape != a.shape else b
                        d = (a.float() - b.float()).abs()
                        if d.mean().item() > (0.6 if k.endswith("bias") else 0.08) * lr * steps:
                            good = False
                    log(f"[{'ok' if good else 'FAIL'}] gpt2-tiny 2-D mesh data=2 x fsdp={world // 2}: losses "
                        f"{[f'{x:.4f}' for x in losses]} shard group {eng.shard_peer_ranks}")
                    ok_all &= good
                    del eng
            except Exception as e:  # noqa: BLE001
                traceback.print_exc()
                log(f"[FAIL] 2-D mesh: {type(e).__name__}: {e}")
                ok_all = False
        # ---- GPT-2 small timing ----
        scfg = gpt2_config("small")
        smodel = GPT2(scfg)
        sparams = smodel.init({"params": rng.PRNGKey(1)}, gb.inputs[:1])["params"]
        stok = torch.randint(0, scfg.vocab_size, (8, 1025), generator=torch.Generator().manual_seed(rank))
        sb = Batch(stok[:, :1024].contiguous().pin_memory(), stok[:, 1:].contiguous().pin_memory())
        timing = () if os.environ.get("B200_SKIP_TIMING") == "1" else \
            (("dp", "dp", "fused"), ("fsdp/fused", "fsdp", "fused"), ("fsdp/gather", "fsdp", "gather"))
        for name, mode, fw in timing:
            try:
                eng = GPT2Engine(scfg, mesh, mode=mode, batch_per_rank=8, seq_len=1024, lr=3e-4, init_params=sparams,
                                 use_graph=True, fsdp_weights=fw)
                s3 = eng.make_state(smodel, rng.PRNGKey(2))
                for _ in range(4):
                    s3, m3 = eng.train_step(s3, None, sb)      # metrics=None: m3 holds this step only
                last_loss = float(m3["loss"][0]) / m3["loss"][1]
                torch.cuda.synchronize()
                evs = []
                for _ in range(10):
                    eng._load_batch(sb)
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(); eng._run_step(); b.record()
                    evs.append((a, b))
                torch.cuda.synchronize()
                ms = sorted(x.elapsed_time(y) for x, y in evs)[len(evs) // 2]
                t = torch.tensor([ms], device=device)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                log(f"gpt2-small {name} x{world}: {t.item():.2f} ms/step, {8 * world / t.item() * 1e3:.0f} samples/s, "
                    f"loss of the last warm-up step {last_loss:.3f}, launches/step {eng.launches_per_step}")
                del eng
                torch.cuda.empty_cache()
            except Exception as e:  # noqa: BLE001
                traceback.print_exc()
                log(f"[FAIL] gpt2-small {name}: {type(e).__name__}: {e}")
                ok_all = False
    log("MULTI-GPU GPT CHECK", "PASS" if ok_all else "FAIL")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok_all else 1)


if __name__ == "__main__":
    main()

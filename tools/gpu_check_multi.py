"""Multi-GPU checks (launch with torchrun, one rank per GPU):

  torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 tools/gpu_check_multi.py

1. fused DP engine (NVLS multicast path, then the P2P path) == generic DP over NCCL;
   replicas bit-identical
2. fused FSDP engine (all-gather fused into the GEMM + reduce-scatter fused with AdamW) == DP engine
3. step timing of each, device-timed, max over ranks
"""
from __future__ import annotations

import os
import sys
import traceback

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from jax_distributed_tuts_b200 import rng
from jax_distributed_tuts_b200.config import dp_config
from jax_distributed_tuts_b200.engine import MLPEngine
from jax_distributed_tuts_b200.launch import init_from_env
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.mlp import Classifier
from jax_distributed_tuts_b200.parallel import dp
from jax_distributed_tuts_b200.state import Batch
from jax_distributed_tuts_b200.tree import tree_leaves, tree_map


def log(*a):
    if dist.get_rank() == 0:
        print(*a, flush=True)


def close(a, b, lr, steps):
    d = (a.float() - b.float().reshape(a.shape)).abs()
    return bool(d.max() <= 2.2 * lr * steps and d.mean() <= 0.06 * lr * steps), d.max().item(), d.mean().item()


def main():
    rank, world, device = init_from_env()
    per_rank, steps, lr = 64, 4, 1e-3
    cfg = dp_config()
    mesh = Mesh((world,), ("data",))
    model = Classifier(cfg.model)
    key = rng.PRNGKey(cfg.seed)
    model_rng, data_rng, label_rng = rng.split(key, 3)
    gB = per_rank * world
    x = torch.randn(gB, 784, generator=rng.uniform_generator(data_rng))
    y = torch.randint(0, 10, (gB,), generator=rng.uniform_generator(label_rng))
    ok_all = True
    with mesh:
        init = dp.init_dp(model_rng, x, model)
        local = dp.shard_batch(Batch(x, y))
        # ---- generic DP over NCCL ----
        to_dev = lambda t: t.to(device)
        gen = init.replace(params=tree_map(to_dev, init.params), rng=init.rng.to(device))
        gen = gen.replace(opt_state=gen.tx.init(gen.params))
        step = dp.make_train_step_dp(mesh, num_minibatches=4)
        lb = local.to(device)
        m_gen = None
        for _ in range(steps):
            gen, m_gen = step(gen, m_gen, lb)
        torch.cuda.synchronize()
        log(f"generic NCCL DP: loss {float(m_gen['loss'][0]) / m_gen['loss'][1]:.5f} count {m_gen['loss'][1]}")
        results = {}
        for name, mode, env in (("dp/nvls", "dp", "0"), ("dp/p2p", "dp", "1"), ("fsdp/nvls", "fsdp", "0"),
                                ("fsdp/p2p", "fsdp", "1")):
            try:
                os.environ["B200DIST_NO_MULTICAST"] = env
                eng = MLPEngine(cfg.model, mesh, mode=mode, batch_per_rank=per_rank, lr=lr, num_minibatches=4,
                                init_params=init.params, min_weight_size=16, use_graph=True)
                st = eng.make_state(model, init.rng)
                pinned = Batch(local.inputs.pin_memory(), local.labels.pin_memory())
                m = None
                for _ in range(steps):
                    st, m = eng.train_step(st, m, pinned)
                torch.cuda.synchronize()
                tree = eng.params_tree()
                good = True
                for k in ("input_dense", "output_dense"):
                    for kk in ("kernel", "bias"):
                        c, mx, mean = close(tree[k][kk], gen.params[k][kk], lr, steps)
                        good &= c
                        if not c:
                            log(f"  [{name}] {k}/{kk} mismatch max {mx:.2e} mean {mean:.2e}")
                # replicas identical
                flat = torch.cat([t.reshape(-1).float() for kk in tree.values() for t in kk.values()])
                ref = flat.clone()
                dist.broadcast(ref, 0)
                same = bool(torch.equal(ref, flat))
                t = torch.tensor([int(same)], device=device)
                dist.all_reduce(t, op=dist.ReduceOp.MIN)
                loss = float(m["loss"][0]) / m["loss"][1]
                lg = float(m_gen["loss"][0]) / m_gen["loss"][1]
                good &= abs(loss - lg) < 2e-2 and m["loss"][1] == m_gen["loss"][1]
                # Adam moments exported in the global per-leaf layout == the generic path's optimizer state
                og = eng.opt_state_global()
                adam = gen.opt_state[0]
                for nm, mine, ref_t in (("mu", og["mu"], adam.mu), ("nu", og["nu"], adam.nu)):
                    for k in ("input_dense", "output_dense"):
                        for kk in ("kernel", "bias"):
                            a, b_ = mine[k][kk].float(), ref_t[k][kk].float().reshape(mine[k][kk].shape)
                            err = ((a - b_).norm() / b_.norm().clamp_min(1e-12)).item()
                            if not err < 8e-2:
                                good = False
                                log(f"  [{name}] {nm} {k}/{kk} rel err {err:.3e}")
                good &= int(og["count"].item()) == steps
                # checkpoint round trip: a fresh engine restored from the file continues like the original
                import tempfile
                from jax_distributed_tuts_b200 import checkpoint as ckpt
                path = os.path.join(tempfile.gettempdir(), f"b200_ckpt_{name.replace('/', '_')}.msgpack")
                ckpt.save_checkpoint(path, st, mesh=mesh)
                dist.barrier()
                eng2 = MLPEngine(cfg.model, mesh, mode=mode, batch_per_rank=per_rank, lr=lr, num_minibatches=4,
                                 init_params=tree_map(torch.zeros_like, init.params), min_weight_size=16, use_graph=False)
                st2 = eng2.make_state(model, init.rng)
                ckpt.restore_engine(eng2, ckpt.load_checkpoint(path))
                st, m = eng.train_step(st, None, pinned)
                st2, m2 = eng2.train_step(st2, None, pinned)
                torch.cuda.synchronize()
                t1, t2 = eng.params_tree(), eng2.params_tree()
                for k in ("input_dense", "output_dense"):
                    for kk in ("kernel", "bias"):
                        d = (t1[k][kk].float() - t2[k][kk].float()).abs().max().item()
                        if not d <= 2e-5:
                            good = False
                            log(f"  [{name}] resume mismatch {k}/{kk}: {d:.3e}")
                if abs(float(m["loss"][0]) - float(m2["loss"][0])) > 1e-2 * max(1.0, abs(float(m["loss"][0]))):
                    good = False
                    log(f"  [{name}] resume loss {float(m2['loss'][0])} vs {float(m['loss'][0])}")
                del eng2
                # timing
                evs = []
                for _ in range(30):
                    eng._load_batch(pinned)
                    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    s.record(); eng._run_step(); e.record()
                    evs.append((s, e))
                torch.cuda.synchronize()
                ms = sorted(a.elapsed_time(b) for a, b in evs)[len(evs) // 2]
                tt = torch.tensor([ms], device=device)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                log(f"[{'ok' if good and t.item() == 1 else 'FAIL'}] {name}: backend={eng.arena.backend} loss {loss:.5f} "
                    f"(generic {lg:.5f}) replicas_identical={bool(t.item())} median step {tt.item() * 1e3:.1f} us "
                    f"launches/step {eng.launches_per_step}")
                ok_all &= good and t.item() == 1
                results[name] = tree
                del eng
            except Exception as e:  # noqa: BLE001
                traceback.print_exc()
                log(f"[FAIL] {name}: {type(e).__name__}: {e}")
                ok_all = False
                break
        # ---- generic tier on the fused collective backend (no NCCL kernels for psum/pmean/gather/scatter) ----
        try:
            from jax_distributed_tuts_b200.comm import collectives as Cc
            Cc.set_backend("fused")
            before = Cc.fused_calls
            gen2 = init.replace(params=tree_map(to_dev, init.params), rng=init.rng.to(device))
            gen2 = gen2.replace(opt_state=gen2.tx.init(gen2.params))
            m2 = None
            for _ in range(steps):
                gen2, m2 = step(gen2, m2, lb)
            torch.cuda.synchronize()
            worst = max((a.float() - b.float()).abs().max().item()
                        for a, b in zip(tree_leaves(gen2.params), tree_leaves(gen.params)))
            good = worst < 2e-5 and abs(float(m2["loss"][0]) - float(m_gen["loss"][0])) < 1e-3 and Cc.fused_calls > before
            # primitives against torch.distributed on odd sizes
            g = torch.Generator(device="cpu").manual_seed(100 + rank)
            for n in (10, 4099, 1 << 20):
                t = torch.randn(n, generator=g).to(device)
                ref = t.clone(); dist.all_reduce(ref)
                got = Cc.psum(t, "data")
                good &= bool((got - ref).abs().max().item() <= 1e-5 * max(1.0, ref.abs().max().item()))
                ga = Cc.all_gather(t.to(torch.bfloat16), "data", axis=0)
                ref_g = torch.empty(world * n, dtype=torch.bfloat16, device=device)
                dist.all_gather_into_tensor(ref_g, t.to(torch.bfloat16))
                good &= bool(torch.equal(ga, ref_g))
            t2 = torch.randn(world * 6, 33, generator=g).to(device)
            rs = Cc.psum_scatter(t2, "data", scatter_dimension=0)
            ref2 = t2.clone(); dist.all_reduce(ref2)
            good &= bool((rs - ref2[rank * 6:(rank + 1) * 6]).abs().max().item() < 1e-4)
            cnt = Cc.psum(torch.tensor([rank + 1], device=device), "data")
            good &= int(cnt.item()) == world * (world + 1) // 2
            log(f"[{'ok' if good else 'FAIL'}] generic DP step + primitives on the fused collective backend: "
                f"max param diff vs NCCL {worst:.2e}, fused calls {Cc.fused_calls - before}")
            ok_all &= good
        except Exception as e:  # noqa: BLE001
            traceback.print_exc()
            log(f"[FAIL] fused collective backend: {type(e).__name__}: {e}")
            ok_all = False
        finally:
            Cc.set_backend("auto")
    log("MULTI-GPU CHECK", "PASS" if ok_all else "FAIL")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok_all else 1)


if __name__ == "__main__":
    main()

"""Headline benchmark: train samples/sec of the data-parallel train step (BASELINE.json metric
"train samples/sec (device-timed, max over ranks) at 1/2/4/8 B200"), on the reference's own
workload by default: MLP 784->512->10 (SiLU, dropout 0.1), bf16 compute / fp32 master params,
AdamW, 16 rows per device in 4 minibatches (global batch 128 on 8 devices; data_paral.py:39-72).

    python bench.py --gpus N --steps K --warmup W            # N = 1: plain python
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0. ``value`` is device-timed (CUDA events around every step, L2
flushed between steps outside the events, sum over K steps, max over ranks); ``e2e`` is the
same step through the public API with the batch copied from pinned host memory and the loss
read back every step, wall-clock, max over ranks.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

CONFIGS = {
    # name: (model, per-gpu batch, minibatches, hidden, input, classes)
    "mlp_dp": dict(model="mlp_784_512_10", per_gpu_batch=16, minibatches=4, hidden=512, inp=784, classes=10,
                   mode="dp"),
    "mlp_dp_large": dict(model="mlp_784_8192_10", per_gpu_batch=32768, minibatches=4, hidden=8192, inp=784,
                         classes=10, mode="dp"),
    "mlp_fsdp": dict(model="mlp_784_512_10", per_gpu_batch=16, minibatches=4, hidden=512, inp=784, classes=10,
                     mode="fsdp"),
    # transformer configurations of BASELINE.json (samples = sequences of 1024 tokens)
    "gpt2s_dp": dict(model="gpt2_small_124M", gpt="small", per_gpu_batch=16, minibatches=1, seq=1024, mode="dp"),
    "gpt2s_fsdp": dict(model="gpt2_small_124M", gpt="small", per_gpu_batch=16, minibatches=1, seq=1024, mode="fsdp"),
    "gpt2m_dp": dict(model="gpt2_medium_355M", gpt="medium", per_gpu_batch=16, minibatches=1, seq=1024, mode="dp"),
    "gpt2m_2d": dict(model="gpt2_medium_355M", gpt="medium", per_gpu_batch=16, minibatches=1, seq=1024, mode="fsdp",
                     mesh2d=True),
    "gpt2_tiny": dict(model="gpt2_tiny", gpt="tiny", per_gpu_batch=4, minibatches=1, seq=128, mode="dp"),
    # 4-stage GPipe over NCCL p2p (BASELINE.json config 4): per_gpu_batch is per *pipeline* (data replica)
    "gpt2s_pp4": dict(model="gpt2_small_124M", gpt="small", per_gpu_batch=8, minibatches=4, seq=1024, mode="pp", stages=4),
}


def reference_arm(args) -> None:
    """The reference cannot be installed or run on this image (recorded in DESIGN.md):
    `pip install --no-index ... /root/reference` fails (no setup.py/pyproject.toml: it is four
    loose scripts) and its imports (jax, flax, optax, ml_collections, absl, termcolor) are not
    available offline."""
    why = "reference is 4 loose JAX scripts: not pip-installable (no setup.py/pyproject) and jax/flax/optax are not in the offline image"
    ref_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "baseline", "_ref")
    try:
        sys.path.insert(0, ref_dir)
        import jax  # noqa: F401
        why = "jax import succeeded unexpectedly but the reference scripts raise NameError at import (util.py:32 uses os without importing it)"
    except Exception:  # noqa: BLE001
        pass
    print(json.dumps({"impl": "reference", "unavailable": why}))


class ClockSampler:
    """Samples SM clock / throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100", "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:  # noqa: BLE001
            self.p = None

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:  # noqa: BLE001
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        clocks, reasons, mx = [], set(), None
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                clocks.append(float(parts[1]))
                mx = float(parts[2])
            except ValueError:
                continue
            for nm, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if clocks:
            clocks.sort()
            # median over the upper half = clocks under load (idle samples sit at the bottom)
            upper = clocks[len(clocks) // 2:]
            out.update(sm_mhz=upper[len(upper) // 2], sm_max_mhz=mx, reasons=sorted(reasons), samples=len(clocks))
        return out


def run_secondary(args, rank, world, device) -> dict:
    """DP GPT-2 small (124M), 16 x 1024 tokens per GPU, 20 timed steps, same timing rules."""
    import torch
    import torch.distributed as dist

    from jax_distributed_tuts_b200 import rng
    from jax_distributed_tuts_b200.engine import GPT2Engine
    from jax_distributed_tuts_b200.mesh import Mesh
    from jax_distributed_tuts_b200.models.gpt2 import GPT2, gpt2_config
    from jax_distributed_tuts_b200.state import Batch

    steps, warm, B, T = 20, 4, 16, 1024
    gcfg = gpt2_config("small")
    model = GPT2(gcfg)
    mesh = Mesh((world,), ("data",))
    g = torch.Generator().manual_seed(1000 + rank)
    tok = torch.randint(0, gcfg.vocab_size, (B, T + 1), generator=g)
    batch = Batch(tok[:, :T].contiguous().pin_memory(), tok[:, 1:].contiguous().pin_memory())
    with mesh:
        params = model.init({"params": rng.PRNGKey(1)}, batch.inputs[:1])["params"]
        eng = GPT2Engine(gcfg, mesh, mode="dp", batch_per_rank=B, seq_len=T, lr=3e-4, init_params=params)
        del params
        st = eng.make_state(model, rng.PRNGKey(2))
        for _ in range(warm):
            st, m = eng.train_step(st, None, batch)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
        evs = []
        for _ in range(steps):
            flush.zero_()
            eng._load_batch(batch)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); eng._run_step(); b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        t = torch.tensor([sum(x.elapsed_time(y) for x, y in evs)], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item() / steps
        toks = B * T * world
        flops = 6.0 * 124.4e6 * toks + 12.0 * gcfg.n_layer * gcfg.d_model * T * toks * 0.5
        launches = int(eng.launches_per_step)
        del eng
        torch.cuda.empty_cache()
    return {"metric": "train samples/sec (device-timed, max over ranks)", "value": B * world / (ms * 1e-3),
            "unit": "samples/s (1 sample = 1024 tokens)", "ms_per_step": ms, "tokens_per_s": toks / (ms * 1e-3),
            "model_tflops_per_gpu": flops / (ms * 1e-3) / 1e12 / world, "steps": steps, "warmup": warm,
            "config": {"model": "gpt2_small_124M", "global_batch": B * world, "seq_len": T, "parallelism": f"dp{world}"},
            "launches_per_step": launches}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "baseline"],
                    help="ours = fused engines; baseline = stock PyTorch+NCCL+cuBLAS reconstruction (baseline/); "
                         "reference = the unmodified reference (not installable here)")
    ap.add_argument("--config", default="mlp_dp", choices=sorted(CONFIGS))
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--per-gpu-batch", type=int, default=0, help="override the config's per-GPU batch (exploration)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--kernel-table", action="store_true",
                    help="after the timed regions, profile 3 more steps with CUPTI and print per-kernel device time to stderr")
    ap.add_argument("--no-extra", action="store_true",
                    help="skip the secondary GPT-2 small measurement that the default config appends as `extra`")
    args = ap.parse_args()
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) == 0:
            reference_arm(args)
        return
    args.warmup = max(args.warmup, 3)

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

    rank, world, device = init_from_env()
    assert device.type == "cuda", "bench.py needs a GPU"
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run"
    c = dict(CONFIGS[args.config])
    if args.per_gpu_batch:
        c["per_gpu_batch"] = args.per_gpu_batch
    cfg = dp_config()
    B = c["per_gpu_batch"]
    is_gpt = "gpt" in c
    if c["mode"] == "pp":
        assert world % c["stages"] == 0, f"{args.config} needs a multiple of {c['stages']} GPUs"
        mesh = Mesh((world // c["stages"], c["stages"]), ("data", "pipe"))
        shard_axis = "data"
    elif c.get("mesh2d") and world >= 4:
        mesh = Mesh((2, world // 2), ("data", "fsdp"))
        shard_axis = "fsdp"
    else:
        mesh = Mesh((world,), ("data",))
        shard_axis = "data"
    key = rng.PRNGKey(cfg.seed)
    model_rng, data_rng, label_rng = rng.split(key, 3)
    # synthetic data of the named shape; a distinct shard per rank, fixed across steps like the
    # reference's single synthetic batch (data_paral.py:113-124)
    gx = rng.uniform_generator(rng.fold_in(data_rng, rank if c["mode"] != "pp" else mesh.axis_index("data")))
    if is_gpt:
        from jax_distributed_tuts_b200.engine import GPT2Engine
        from jax_distributed_tuts_b200.models.gpt2 import GPT2, gpt2_config
        gcfg = gpt2_config(c["gpt"])
        model = GPT2(gcfg)
        T = c["seq"]
        tok = torch.randint(0, gcfg.vocab_size, (B, T + 1), generator=gx)
        x, y = tok[:, :T].contiguous().pin_memory(), tok[:, 1:].contiguous().pin_memory()
        h2d_bytes = int(x.numel() * 8 + y.numel() * 8)
    else:
        cfg.model.hidden_size = c["hidden"]
        model = Classifier(cfg.model)
        x = torch.randn(B, c["inp"], generator=gx).pin_memory()
        y = torch.randint(0, c["classes"], (B,), generator=rng.uniform_generator(rng.fold_in(label_rng, rank))).pin_memory()
        h2d_bytes = int(x.numel() * 4 + y.numel() * 8)
    with mesh:
        if args.impl == "baseline":
            from baseline.torch_nccl import GPT2 as TGPT2, MLP as TMLP, BaselineTrainer
            torch.manual_seed(cfg.seed)
            if is_gpt:
                tm = TGPT2(gcfg.n_layer, gcfg.n_head, gcfg.d_model, 50304, gcfg.n_ctx).to(device)
                lr0 = 3e-4
            else:
                tm = TMLP(c["inp"], c["hidden"], c["classes"], 0.1).to(device)
                lr0 = cfg.optimizer.learning_rate
            tm.train()
            trainer = BaselineTrainer(tm, mode=c["mode"], lr=lr0, minibatches=c["minibatches"], world=world, rank=rank)
            xd, yd = torch.empty_like(x, device=device), torch.empty_like(y, device=device)
            last = {}

            def load_batch():
                xd.copy_(x, non_blocking=True)
                yd.copy_(y, non_blocking=True)

            def run_step():
                last["loss"] = trainer.step(xd, yd)

            def api_step():
                load_batch()
                run_step()
                return last["loss"], 1.0

            backend, launches_per_step = "nccl" if world > 1 else "local", 0
        elif c["mode"] == "pp":
            from jax_distributed_tuts_b200.parallel import pp
            from jax_distributed_tuts_b200.tree import tree_map
            stage = pp.GPT2Stage(model, mesh.axis_index("pipe"), c["stages"])
            st0 = pp.init_pp(model_rng, x[:1], stage, learning_rate=3e-4)
            st0 = st0.replace(params=tree_map(lambda t: t.to(device), st0.params), rng=st0.rng.to(device))
            st0 = st0.replace(opt_state=st0.tx.init(st0.params))
            pp_step = pp.make_train_step_pp(mesh, stage_model=stage, num_microbatches=c["minibatches"], pipe_axis="pipe",
                                            data_axis="data" if mesh.axis_size("data") > 1 else None)
            xd, yd = torch.empty_like(x, device=device), torch.empty_like(y, device=device)
            holder = {"state": st0, "metrics": None}

            def load_batch():
                xd.copy_(x, non_blocking=True)
                yd.copy_(y, non_blocking=True)

            def run_step():
                holder["state"], holder["metrics"] = pp_step(holder["state"], None, Batch(xd, yd))

            def api_step():
                load_batch()
                run_step()
                return holder["metrics"]["loss"][0], float(holder["metrics"]["loss"][1])

            backend, launches_per_step = "nccl-p2p", 0
        else:
            if is_gpt:
                params = model.init({"params": model_rng}, x[:1])["params"]      # random init of the named architecture
                eng = GPT2Engine(gcfg, mesh, mode=c["mode"], batch_per_rank=B, seq_len=T, lr=3e-4,
                                 num_minibatches=c["minibatches"], init_params=params, use_graph=not args.no_graph,
                                 shard_axis=shard_axis)
                del params
                state0 = eng.make_state(model, rng.rebase(model_rng))
            else:
                init_state = dp.init_dp(model_rng, x[:4], model)
                eng = MLPEngine(cfg.model, mesh, mode=c["mode"], batch_per_rank=B, lr=cfg.optimizer.learning_rate,
                                num_minibatches=c["minibatches"], init_params=init_state.params, input_size=c["inp"],
                                use_graph=not args.no_graph,
                                min_weight_size=16 if c["mode"] == "fsdp" else 2 ** 18)
                state0 = eng.make_state(model, init_state.rng)
            pub_step = dp.make_train_step_dp(mesh)      # the public entry point: step(state, metrics, batch)
            batch = Batch(x, y)
            holder = {"state": state0, "metrics": None}

            def load_batch():
                eng._load_batch(batch)

            def run_step():
                eng._run_step()

            def api_step():
                holder["state"], holder["metrics"] = pub_step(holder["state"], None, batch)
                return holder["metrics"]["loss"][0], 1.0

        def sync():
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
                torch.cuda.synchronize()

        # ---- warm-up (untimed): includes the eager step and the graph capture ----
        for _ in range(args.warmup):
            api_step()
        sync()
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)  # > 126 MB L2

        # ---- device-timed region: events around each step, L2 flushed in between ----
        sampler = ClockSampler(torch.cuda.current_device()) if rank == 0 else None
        evs = []
        sync()
        for _ in range(args.steps):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            load_batch()                # H2D of the step's inputs is outside the events (kernel-only number)
            s.record()
            run_step()
            e.record()
            evs.append((s, e))
        sync()
        dev_ms = sum(s.elapsed_time(e) for s, e in evs)
        clocks = sampler.stop() if sampler else None

        # ---- end-to-end region: public API, pinned H2D every step, loss read back every step ----
        # The loss of every step is copied to pinned host memory on the stream (4 bytes, counted in
        # d2h_bytes_per_step) and consumed by the host one step later, so the read does not drain
        # the launch queue; the last value is read after the loop.
        e2e_ms, loss_val = None, None
        if not args.no_e2e:
            pins = [torch.empty(1, dtype=torch.float32).pin_memory() for _ in range(2)]
            done = [torch.cuda.Event(), torch.cuda.Event()]
            pending, pending_den = None, 1.0
            sync()
            t0 = time.perf_counter()
            for i in range(args.steps):
                lt, den = api_step()
                slot = i & 1
                pins[slot].copy_(lt.detach().reshape(1).float(), non_blocking=True)
                done[slot].record()
                if pending is not None:
                    done[pending].synchronize()
                    loss_val = pins[pending].item() / pending_den
                pending, pending_den = slot, den
            done[pending].synchronize()
            loss_val = pins[pending].item() / pending_den
            sync()
            e2e_ms = (time.perf_counter() - t0) * 1e3

        if args.kernel_table and rank == 0:     # diagnostic only: runs after every reported number was taken
            from torch.profiler import ProfilerActivity, profile
            sync_local = torch.cuda.synchronize
            sync_local()
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                for _ in range(3):
                    run_step()
                sync_local()
            rows_ = sorted(prof.key_averages(), key=lambda ev: -ev.device_time_total)
            tot = sum(ev.device_time_total for ev in rows_) / 3.0
            print(f"[kernel table] {tot / 1e3:.3f} ms of kernel time per step (sum over kernels, PDL overlap counted twice)", file=sys.stderr)
            for ev in rows_[:28]:
                print(f"  {ev.device_time_total / 3e3:8.3f} ms  x{ev.count // 3:4d}  {ev.key[:110]}", file=sys.stderr)

        t = torch.tensor([dev_ms, e2e_ms or 0.0], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_max = t[0].item(), t[1].item()
        if args.impl == "baseline" or c["mode"] == "pp":
            final_loss = loss_val
        else:
            rows = eng._rows_per_step()
            final_loss = (loss_val / rows) if loss_val is not None else None
            backend, launches_per_step = eng.arena.backend, int(getattr(eng, "launches_per_step", 0))

    if rank == 0:
        gb = B * world if c["mode"] != "pp" else B * (world // c["stages"])
        ms_per_step = dev_ms / args.steps
        out = {
            "metric": "train samples/sec (device-timed, max over ranks)",
            "value": gb / (ms_per_step * 1e-3),
            "unit": "samples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic (random-normal inputs, uniform labels, random-init weights)",
            "impl": args.impl,
            "config": {"model": c["model"], "global_batch": gb, "per_gpu_batch": B, "minibatches": c["minibatches"],
                       "seq_len": c.get("seq"), "parallelism": ("dp2xfsdp%d" % (world // 2)) if (c.get("mesh2d") and world >= 4) else
                                      (f"dp{world // c['stages']}xpp{c['stages']}" if c["mode"] == "pp" else f"{c['mode']}{world}"),
                       "optimizer": "adamw",
                       "name": args.config, "l2": "flushed between timed steps (256 MiB write)",
                       "cuda_graph": (not args.no_graph) and args.impl == "ours", "comm": backend},
            "gpu_launches": launches_per_step * args.steps,
            "launches_per_step": launches_per_step,
            "clocks": clocks,
            "final_loss": final_loss,
        }
        if e2e_ms is not None:
            out["e2e"] = {"value": gb / (e2e_max / args.steps * 1e-3), "unit": "samples/s",
                          "ms_per_step": e2e_max / args.steps,
                          "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 4,
                          "readback": "loss copied to pinned host memory every step, consumed by the host one step later"}
        if is_gpt:
            toks = gb * c["seq"]
            out["tokens_per_s"] = toks / (ms_per_step * 1e-3)
            n_params = 124.4e6 if c["gpt"] == "small" else 354.8e6 if c["gpt"] == "medium" else 1e6
            flops = 6.0 * n_params * toks + 12.0 * gcfg.n_layer * gcfg.d_model * c["seq"] * toks * 0.5
            out["model_tflops_per_gpu"] = flops / (ms_per_step * 1e-3) / 1e12 / world
    want_extra = (args.config == "mlp_dp" and args.impl == "ours" and not args.no_extra)
    extra = None
    if want_extra:
        # secondary measurement, same protocol, transformer configuration of BASELINE.json (DP GPT-2 small):
        # release this run's engine first, then run the GPT-2 config in the same processes
        try:
            del eng, holder, state0
            torch.cuda.empty_cache()
            if world > 1:
                dist.barrier()
            extra = run_secondary(args, rank, world, device)
        except Exception as e:  # noqa: BLE001  (never lose the headline line because of the extra)
            extra = {"error": f"{type(e).__name__}: {e}"[:200]}
    if rank == 0:
        if extra is not None:
            out["extra"] = {"gpt2s_dp": extra}
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""Headline benchmark: train samples/sec of the data-parallel train step (BASELINE.json metric
"train samples/sec (device-timed, max over ranks) at 1/2/4/8 B200"), on the reference's own
workload by default: MLP 784->512->10 (SiLU, dropout 0.1), bf16 compute / fp32 master params,
AdamW, 16 rows per device in 4 minibatches (global batch 128 on 8 devices; data_paral.py:39-72).

    python bench.py --gpus N --steps K --warmup W            # N = 1: plain python
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0. ``value`` is device-timed (CUDA events around every step, L2
flushed between steps outside the events, sum over K steps, max over ranks); ``e2e`` is the
same step through the public API with the batch copied from pinned host memory and the loss
read back every step, wall-clock, max over ranks. Next to the headline:

    step_ms        min / median / p90 / max of the per-step device times (max over ranks per step)
    selfcheck      N > 1: one step of the fused update against NCCL all-reduce + fp32 PyTorch AdamW on
                   the same gradients, and bit-identity of the replicas, before anything is timed
    exposed_comm   N > 1: step time minus the same rank's step with every cross-GPU access stubbed
    roofline       the slower of FLOPs at the measured GEMM peak and NVLink bytes at the measured link
                   rate, divided by the measured step time
    extra          the other BASELINE.json configurations (GPT-2 small DP / FSDP, GPT-2 medium 2-D mesh,
                   4-stage pipeline from 4 GPUs), same protocol with fewer steps
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


SUSTAINED_TFLOPS, NVLINK_GBS = 1406.5, 770.0      # MEASURED_PEAKS.json (sustained bf16 GEMM); measured peer copy, per direction
N_PARAMS = {"small": 124.4e6, "medium": 354.8e6, "tiny": 1e6}


def model_flops_per_sample(c, gcfg=None) -> float:
    """fwd+bwd FLOPs of one sample (GPT: one sequence; 6 x params x tokens + causal attention)."""
    if "gpt" in c:
        T = c["seq"]
        return 6.0 * N_PARAMS[c["gpt"]] * T + 12.0 * gcfg.n_layer * gcfg.d_model * T * T * 0.5
    return 6.0 * (c["inp"] * c["hidden"] + c["hidden"] * c["classes"])


def nvlink_bytes_per_rank(c, n_params: float, world: int, mesh_shape) -> float:
    """Bytes that must cross one GPU's NVLink in the busier direction per step (algorithmic minimum)."""
    if world == 1:
        return 0.0
    f = (world - 1) / world
    if c["mode"] == "dp":          # reduce: every rank serves its copy of the other owners' slices (fp32)
        return 4.0 * n_params * f
    if c["mode"] == "fsdp":        # weights gathered for forward and backward (bf16) vs fp32 gradient reduce egress
        S = mesh_shape[-1]
        return max(4.0 * n_params * f, 2 * 2.0 * n_params * (S - 1) / S)
    return 0.0                     # pipeline: activations only (reported separately)


def build_runner(name: str, args, rank: int, world: int, device, impl: str):
    """Everything one configuration needs: engine / trainer, pinned synthetic batch, the callables
    ``load_batch`` (H2D of the step's inputs), ``run_step`` (the step's launches), ``api_step``
    (the public entry point: H2D + step, returns (loss tensor, denominator))."""
    import torch

    from jax_distributed_tuts_b200 import rng
    from jax_distributed_tuts_b200.config import dp_config
    from jax_distributed_tuts_b200.mesh import Mesh
    from jax_distributed_tuts_b200.parallel import dp
    from jax_distributed_tuts_b200.state import Batch

    c = dict(CONFIGS[name])
    if args.per_gpu_batch and name == args.config:
        c["per_gpu_batch"] = args.per_gpu_batch
    cfg = dp_config()
    B = c["per_gpu_batch"]
    is_gpt = "gpt" in c
    if c["mode"] == "pp":
        assert world % c["stages"] == 0, f"{name} needs a multiple of {c['stages']} GPUs"
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
    gcfg = None
    if is_gpt:
        from jax_distributed_tuts_b200.models.gpt2 import GPT2, gpt2_config
        gcfg = gpt2_config(c["gpt"])
        model = GPT2(gcfg)
        T = c["seq"]
        tok = torch.randint(0, gcfg.vocab_size, (B, T + 1), generator=gx)
        x, y = tok[:, :T].contiguous().pin_memory(), tok[:, 1:].contiguous().pin_memory()
        h2d_bytes = int(x.numel() * 8 + y.numel() * 8)
    else:
        from jax_distributed_tuts_b200.models.mlp import Classifier
        cfg.model.hidden_size = c["hidden"]
        model = Classifier(cfg.model)
        x = torch.randn(B, c["inp"], generator=gx).pin_memory()
        y = torch.randint(0, c["classes"], (B,), generator=rng.uniform_generator(rng.fold_in(label_rng, rank))).pin_memory()
        h2d_bytes = int(x.numel() * 4 + y.numel() * 8)
    R = dict(name=name, c=c, mesh=mesh, gcfg=gcfg, B=B, h2d_bytes=h2d_bytes, eng=None, backend="local", is_gpt=is_gpt)
    R["global_batch"] = B * world if c["mode"] != "pp" else B * (world // c["stages"])
    R["rows_den"] = 1.0
    with mesh:
        if impl == "baseline":
            from baseline.torch_nccl import GPT2 as TGPT2, MLP as TMLP, BaselineTrainer
            torch.manual_seed(cfg.seed)
            if is_gpt:
                tm = TGPT2(gcfg.n_layer, gcfg.n_head, gcfg.d_model, 50304, gcfg.n_ctx).to(device)
                lr0 = 3e-4
            else:
                tm = TMLP(c["inp"], c["hidden"], c["classes"], 0.1).to(device)
                lr0 = cfg.optimizer.learning_rate
            tm.train()
            trainer = BaselineTrainer(tm, mode=c["mode"], lr=lr0, minibatches=c["minibatches"], world=world, rank=rank,
                                      stages=c.get("stages", 1), graph=not args.no_graph)
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

            R.update(backend="nccl" if world > 1 else "local", trainer=trainer)
        elif c["mode"] == "pp":
            from jax_distributed_tuts_b200.engine.pp import PipelineEngine
            params = model.init({"params": model_rng}, x[:1])["params"]
            eng = PipelineEngine(gcfg, mesh, batch_per_replica=B, seq_len=c["seq"], num_microbatches=c["minibatches"], lr=3e-4,
                                 init_params=params, use_graph=not args.no_graph)
            del params
            batch = Batch(x, y)

            def load_batch():
                eng.load_batch(batch)

            def run_step():
                eng.run_step()

            def api_step():
                loss_sum, count = eng.train_step(batch)
                return loss_sum, float(count)

            R.update(eng=eng, backend="nccl-p2p")
        else:
            from jax_distributed_tuts_b200.engine import GPT2Engine, MLPEngine
            if is_gpt:
                params = model.init({"params": model_rng}, x[:1])["params"]      # random init of the named architecture
                eng = GPT2Engine(gcfg, mesh, mode=c["mode"], batch_per_rank=B, seq_len=c["seq"], lr=3e-4,
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

            R.update(eng=eng, backend=eng.arena.backend, holder=holder)
            R["rows_den"] = float(eng._rows_per_step())
    R.update(load_batch=load_batch, run_step=run_step, api_step=api_step)
    return R


def _sync(world):
    import torch
    import torch.distributed as dist
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def time_steps(R, steps: int, world: int, flush, device):
    """K steps, CUDA events around each, L2 flushed in between (outside the events). Returns the
    per-step times (max over ranks per step) and the sum over steps (max over ranks of the sums)."""
    import torch
    import torch.distributed as dist
    evs = []
    _sync(world)
    for _ in range(steps):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        R["load_batch"]()           # H2D of the step's inputs is outside the events (kernel-only number)
        s.record()
        R["run_step"]()
        e.record()
        evs.append((s, e))
    _sync(world)
    t = torch.tensor([s.elapsed_time(e) for s, e in evs], dtype=torch.float64, device=device)
    tot = t.sum().reshape(1)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    return t.cpu().tolist(), float(tot.item())


def _stats(per_step):
    xs = sorted(per_step)
    n = len(xs)
    return {"min": xs[0], "median": xs[n // 2], "p90": xs[min(n - 1, int(0.9 * n))], "max": xs[-1]}


def measure(R, args, steps: int, warmup: int, rank: int, world: int, device, flush, *, e2e: bool, clocks: bool,
            selfcheck: bool, exposed: bool) -> dict:
    """The whole protocol for one configuration; returns the JSON-able record (valid on rank 0)."""
    import torch
    import torch.distributed as dist
    c, eng = R["c"], R["eng"]
    out = {}
    fused = eng is not None and hasattr(eng, "selfcheck")
    # ---- correctness first: N > 1, the first real step against an independent path ----
    if selfcheck and fused and world > 1:
        try:
            out["selfcheck"] = eng.selfcheck(lambda: R["api_step"]())
        except Exception as e:  # noqa: BLE001
            out["selfcheck"] = {"ok": False, "error": f"{type(e).__name__}: {e}"[:200]}
    # ---- warm-up (untimed): includes the eager step and the graph capture ----
    for _ in range(warmup):
        R["api_step"]()
    _sync(world)
    sampler = ClockSampler(torch.cuda.current_device()) if (clocks and rank == 0) else None
    per_step, dev_ms = time_steps(R, steps, world, flush, device)
    ms_per_step = dev_ms / steps
    gb = R["global_batch"]
    # ---- end-to-end region: public API, pinned H2D every step, loss read back every step ----
    # The loss of every step is copied to pinned host memory on the stream (4 bytes, counted in
    # d2h_bytes_per_step) and consumed by the host one step later, so the read does not drain
    # the launch queue; the last value is read after the loop.
    loss_val, e2e_rec = None, None
    if e2e:
        pins = [torch.empty(1, dtype=torch.float32).pin_memory() for _ in range(2)]
        done = [torch.cuda.Event(), torch.cuda.Event()]
        pending, pending_den = None, 1.0
        _sync(world)
        t0 = time.perf_counter()
        for i in range(steps):
            lt, den = R["api_step"]()
            slot = i & 1
            pins[slot].copy_(lt.detach().reshape(1).float(), non_blocking=True)
            done[slot].record()
            if pending is not None:
                done[pending].synchronize()
                loss_val = pins[pending].item() / pending_den
            pending, pending_den = slot, den
        done[pending].synchronize()
        loss_val = pins[pending].item() / pending_den
        _sync(world)
        t = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = t.item() / steps
        e2e_rec = {"value": gb / (e2e_ms * 1e-3), "unit": "samples/s", "ms_per_step": e2e_ms,
                   "h2d_bytes_per_step": R["h2d_bytes"], "d2h_bytes_per_step": 4,
                   "readback": "loss copied to pinned host memory every step, consumed by the host one step later"}
        if loss_val is not None:
            loss_val = loss_val / R["rows_den"]
    # (the clock sampler covers both timed regions: the device-timed one alone lasts tens of milliseconds on
    #  the MLP, less than nvidia-smi's sampling period)
    clk = sampler.stop() if sampler else None
    # ---- exposed communication: the same step with every cross-GPU access stubbed out ----
    if exposed and fused and world > 1:
        try:
            nstub = max(5, steps // 4)
            eng.set_stub_collectives(True)
            for _ in range(3):
                R["load_batch"](); R["run_step"]()
            _, stub_ms = time_steps(R, nstub, world, flush, device)
            stub_ms /= nstub
            eng.set_stub_collectives(False)
            for _ in range(3):
                R["load_batch"](); R["run_step"]()
            _sync(world)
            out["exposed_comm_ms"] = ms_per_step - stub_ms
            out["ms_per_step_comm_stubbed"] = stub_ms
        except Exception as e:  # noqa: BLE001
            out["exposed_comm_ms"] = None
            out["exposed_comm_error"] = f"{type(e).__name__}: {e}"[:200]
    launches = int(getattr(eng, "launches_per_step", 0)) if eng is not None else 0
    par = ("dp2xfsdp%d" % (world // 2)) if (c.get("mesh2d") and world >= 4) else \
        (f"dp{world // c['stages']}xpp{c['stages']}" if c["mode"] == "pp" else f"{c['mode']}{world}")
    out.update({
        "metric": "train samples/sec (device-timed, max over ranks)",
        "value": gb / (ms_per_step * 1e-3), "unit": "samples/s",
        "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms_per_step, "step_ms": _stats(per_step),
        "config": {"model": c["model"], "global_batch": gb, "per_gpu_batch": R["B"], "minibatches": c["minibatches"],
                   "seq_len": c.get("seq"), "parallelism": par, "optimizer": "adamw", "name": R["name"],
                   "l2": "flushed between timed steps (256 MiB write)",
                   "cuda_graph": (not args.no_graph), "comm": R["backend"]},
        "gpu_launches": launches * steps, "launches_per_step": launches,
        "final_loss": loss_val,
    })
    if clk is not None:
        out["clocks"] = clk
    if e2e_rec is not None:
        out["e2e"] = e2e_rec
    # ---- roofline: the slower of compute at the measured peak and NVLink bytes at the measured rate ----
    n_params = N_PARAMS.get(c.get("gpt", ""), 407050.0)
    flops = model_flops_per_sample(c, R["gcfg"]) * gb / world                      # per GPU per step
    nv = nvlink_bytes_per_rank(c, n_params, world, R["mesh"].shape)
    t_flops, t_nv = flops / (SUSTAINED_TFLOPS * 1e12) * 1e3, nv / (NVLINK_GBS * 1e9) * 1e3
    out["roofline"] = {"flops_ms_at_measured_peak": t_flops, "nvlink_ms_at_measured_rate": t_nv,
                       "nvlink_bytes_per_gpu_per_step": nv, "bound_ms": max(t_flops, t_nv),
                       "frac": max(t_flops, t_nv) / ms_per_step,
                       "peaks": f"{SUSTAINED_TFLOPS} TFLOP/s bf16 sustained (MEASURED_PEAKS.json), {NVLINK_GBS} GB/s per direction"}
    if R["is_gpt"]:
        out["tokens_per_s"] = gb * c["seq"] / (ms_per_step * 1e-3)
        out["model_tflops_per_gpu"] = flops / (ms_per_step * 1e-3) / 1e12
    if c["mode"] == "pp" and eng is not None and hasattr(eng, "report"):
        out["pipeline"] = eng.report()
    return out


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
                    help="skip the other BASELINE.json configurations that the default config appends as `extra`")
    ap.add_argument("--extra-budget-s", type=float, default=330.0,
                    help="wall-clock budget for the `extra` configurations (those that do not fit are reported as skipped)")
    args = ap.parse_args()
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) == 0:
            reference_arm(args)
        return
    args.warmup = max(args.warmup, 3)
    t_start = time.perf_counter()

    import torch
    import torch.distributed as dist

    from jax_distributed_tuts_b200.launch import init_from_env

    rank, world, device = init_from_env()
    assert device.type == "cuda", "bench.py needs a GPU"
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)  # > 126 MB L2

    R = build_runner(args.config, args, rank, world, device, args.impl)
    with R["mesh"]:
        out = measure(R, args, args.steps, args.warmup, rank, world, device, flush, e2e=not args.no_e2e, clocks=True,
                      selfcheck=args.impl == "ours", exposed=args.impl == "ours")
        if args.kernel_table and rank == 0:     # diagnostic only: runs after every reported number was taken
            from torch.profiler import ProfilerActivity, profile
            torch.cuda.synchronize()
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                for _ in range(3):
                    R["run_step"]()
                torch.cuda.synchronize()
            rows_ = sorted(prof.key_averages(), key=lambda ev: -ev.device_time_total)
            tot = sum(ev.device_time_total for ev in rows_) / 3.0
            print(f"[kernel table] {tot / 1e3:.3f} ms of kernel time per step (sum over kernels, PDL overlap counted twice)", file=sys.stderr)
            for ev in rows_[:28]:
                print(f"  {ev.device_time_total / 3e3:8.3f} ms  x{ev.count // 3:4d}  {ev.key[:110]}", file=sys.stderr)
    out.update({"higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic (random-normal inputs / uniform tokens, uniform labels, random-init weights)",
                "impl": args.impl})
    # ---- the other BASELINE.json configurations, same protocol, fewer steps ----
    if args.config == "mlp_dp" and args.impl == "ours" and not args.no_extra:
        names = ["gpt2s_dp", "gpt2s_fsdp", "mlp_fsdp", "gpt2m_2d"] + (["gpt2s_pp4"] if world % 4 == 0 else [])
        extra = {}
        # The headline above is measured; an extra that hangs (a lost peer, a wedged collective) must not
        # take it along. Past the deadline every rank's watchdog ends its process, rank 0 after printing
        # the line with what it has.
        import threading
        deadline = t_start + args.extra_budget_s + 150.0
        finished = threading.Event()

        def watchdog():
            while not finished.wait(1.0):
                if time.perf_counter() > deadline:
                    if rank == 0:
                        out["extra"] = dict(extra, watchdog="an extra configuration did not finish in time; the process was ended")
                        print(json.dumps(out), flush=True)
                    os._exit(0)

        threading.Thread(target=watchdog, daemon=True).start()
        R.clear()
        del R
        for nm in names:
            torch.cuda.empty_cache()
            if world > 1:
                dist.barrier()
            spent = torch.tensor([time.perf_counter() - t_start], device=device)
            if world > 1:
                dist.all_reduce(spent, op=dist.ReduceOp.MAX)       # every rank takes the same decision
            if spent.item() > args.extra_budget_s:
                extra[nm] = {"skipped": f"time budget of {args.extra_budget_s:.0f} s used up"}
                continue
            try:
                Rx = build_runner(nm, args, rank, world, device, "ours")
                with Rx["mesh"]:
                    st, wu = (100, 5) if nm.startswith("mlp") else (12, 3)
                    extra[nm] = measure(Rx, args, st, wu, rank, world, device, flush, e2e=False, clocks=False,
                                        selfcheck=True, exposed=nm in ("gpt2s_dp", "gpt2s_fsdp", "mlp_fsdp"))
                Rx.clear()
                del Rx
            except Exception as e:  # noqa: BLE001  (never lose the headline line because of an extra)
                extra[nm] = {"error": f"{type(e).__name__}: {e}"[:300]}
        finished.set()
        out["extra"] = extra
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

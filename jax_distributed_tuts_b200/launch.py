"""Process bootstrap: one process per GPU under torchrun, or N CPU processes over gloo.

``sim_multiCPU_dev`` is the counterpart of the reference's fake-cluster switch
(util.py:31-38: ``--xla_force_host_platform_device_count=N`` + ``CUDA_VISIBLE_DEVICES=""``):
it runs a function on N local CPU processes connected with gloo so that every mesh /
collective / strategy code path executes without a GPU.
"""
from __future__ import annotations

import os
import socket
import tempfile
import traceback
from typing import Any, Callable, Optional

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def init_from_env(backend: Optional[str] = None) -> tuple[int, int, torch.device]:
    """Initialises torch.distributed from RANK/WORLD_SIZE/LOCAL_RANK/MASTER_* (torchrun).

    Sets the CUDA device *before* any context is created (the reference only
    works because JAX creates its backend lazily). Returns (rank, world_size, device).
    """
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    # More hardware launch queues than the default 8 (read when the CUDA context is created): a step uses a
    # compute stream, side streams and one NCCL stream per communicator, and two streams that share a queue
    # pick up each other's waits.
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
    use_cuda = torch.cuda.is_available() and os.environ.get("B200DIST_FORCE_CPU", "0") != "1"
    if use_cuda:
        torch.cuda.set_device(local % torch.cuda.device_count())
        device = torch.device("cuda", torch.cuda.current_device())
    else:
        device = torch.device("cpu")
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = backend or ("nccl" if use_cuda else "gloo")
        kw = {}
        if backend == "nccl":
            kw["device_id"] = device
            # Failure detection: a collective whose peer died must abort the job, not hang.
            # NCCL's watchdog tears the communicator down after the timeout; our own kernels that wait for
            # a peer trap after B200DIST_COMM_TIMEOUT_MS (csrc/comm.cuh), which surfaces as a CUDA error.
            os.environ.setdefault("TORCH_NCCL_ASYNC_ERROR_HANDLING", "1")
            os.environ.setdefault("TORCH_NCCL_BLOCKING_WAIT", "0")
            import datetime
            kw["timeout"] = datetime.timedelta(seconds=int(os.environ.get("B200DIST_NCCL_TIMEOUT_S", "600")))
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, world, device


def _worker(rank: int, world: int, port: int, fn: Callable, args: tuple, result_dir: str,
            timeout_s: Optional[float] = None) -> None:
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), B200DIST_FORCE_CPU="1")
    torch.set_num_threads(1)
    try:
        kw = {}
        if timeout_s is not None:      # a peer that dies must fail the others' collectives after this long
            import datetime
            kw["timeout"] = datetime.timedelta(seconds=timeout_s)
        dist.init_process_group("gloo", rank=rank, world_size=world,
                                init_method=f"tcp://127.0.0.1:{port}", **kw)
        out = fn(rank, world, *args)
        torch.save(out, os.path.join(result_dir, f"rank{rank}.pt"))
        dist.barrier()
    except Exception:  # noqa: BLE001
        with open(os.path.join(result_dir, f"rank{rank}.err"), "w") as f:
            f.write(traceback.format_exc())
        raise
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def sim_multiCPU_dev(fn: Callable[..., Any], device_count: int = 8, args: tuple = (),
                     timeout_s: Optional[float] = None) -> list:
    """Runs ``fn(rank, world, *args)`` on ``device_count`` CPU processes (gloo) and returns the
    per-rank results. The fake-cluster mode of the reference (util.py:31-38). ``timeout_s`` bounds
    every collective: when a rank dies, the survivors' next collective raises after that long and
    the call fails with the workers' tracebacks instead of hanging (failure detection)."""
    port = _free_port()
    with tempfile.TemporaryDirectory() as d:
        ctx = mp.get_context("spawn")
        procs = [ctx.Process(target=_worker, args=(r, device_count, port, fn, args, d, timeout_s))
                 for r in range(device_count)]
        for p in procs:
            p.start()
        for p in procs:
            p.join()
        errs = []
        for r in range(device_count):
            ep = os.path.join(d, f"rank{r}.err")
            if os.path.exists(ep):
                errs.append(f"--- rank {r} ---\n" + open(ep).read())
        if errs or any(p.exitcode != 0 for p in procs):
            raise RuntimeError("sim_multiCPU_dev worker failed:\n" + "\n".join(errs))
        return [torch.load(os.path.join(d, f"rank{r}.pt"), weights_only=False) for r in range(device_count)]


def _run_script(rank: int, world: int, script: str, argv: list) -> None:
    import runpy
    import sys
    sys.argv = [script] + list(argv)
    runpy.run_path(script, run_name="__main__")


def main(argv=None) -> None:
    """``python -m jax_distributed_tuts_b200.launch --cpu N script.py [args]``: runs a script on N CPU
    processes joined by gloo (RANK / WORLD_SIZE / MASTER_* set like torchrun does), the analogue of
    the reference's ``sim_multiCPU_dev()`` call at the top of every script. For GPUs use torchrun."""
    import argparse
    ap = argparse.ArgumentParser(prog="jax_distributed_tuts_b200.launch")
    ap.add_argument("--cpu", type=int, default=8, help="number of simulated devices (CPU processes)")
    ap.add_argument("script")
    ap.add_argument("args", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    sim_multiCPU_dev(_run_script, device_count=a.cpu, args=(os.path.abspath(a.script), a.args))


if __name__ == "__main__":
    main()

"""Pipeline parallelism: what the reference's pipeline_parallel.py stub was heading for, per
BASELINE.json - a GPipe schedule over 4 stages with microbatched stage-to-stage transfers.
A small GPT-2 (8 layers) is split over the "pipe" axis of a (data, pipe) mesh.

On CPU (the fake cluster) and with ``B200DIST_ENGINE=0`` this is the generic tier (autograd, eager
launches, ``parallel/pp.py``); on GPUs under torchrun it is the fused pipeline engine (``engine/pp.py``:
per-stage flat buffers, a CUDA graph per microbatch, NCCL p2p on its own streams)."""
import os

from _common import launch

import torch

from jax_distributed_tuts_b200 import rng
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.gpt2 import GPT2, gpt2_config
from jax_distributed_tuts_b200.parallel import dp, pp
from jax_distributed_tuts_b200.state import Batch
from jax_distributed_tuts_b200.utils import print_metrics

N_STAGES, MICROBATCHES = 4, 4


def worker(rank, world):
    n_data = max(1, world // N_STAGES)
    n_stages = world // n_data
    mesh = Mesh((n_data, n_stages), ("data", "pipe"))
    cfg = gpt2_config("tiny", n_layer=8)
    model = GPT2(cfg)
    g = torch.Generator().manual_seed(0)
    tok = torch.randint(0, cfg.vocab_size, (8 * n_data, 129), generator=g)
    batch = Batch(tok[:, :128].contiguous(), tok[:, 1:].contiguous())
    if mesh.device.type == "cuda" and os.environ.get("B200DIST_ENGINE", "1") == "1":
        from jax_distributed_tuts_b200.engine.pp import PipelineEngine
        d = mesh.axis_index("data")
        local = Batch(batch.inputs[d * 8:(d + 1) * 8].pin_memory(), batch.labels[d * 8:(d + 1) * 8].pin_memory())
        with mesh:
            params = model.init({"params": rng.PRNGKey(0)}, local.inputs[:1])["params"]
            eng = PipelineEngine(cfg, mesh, batch_per_replica=8, seq_len=128, num_microbatches=MICROBATCHES, lr=1e-3,
                                 init_params=params)
            for _ in range(11):
                loss_sum, count = eng.train_step(local)
            if rank == 0:
                print(f"mesh data={n_data} x pipe={n_stages}, layers per stage: {[len(r) for r in eng.split]} ({eng.schedule})")
                print_metrics({"loss": (loss_sum, count)}, "pp - fused pipeline engine")
        return
    with mesh:
        stage = pp.GPT2Stage(model, mesh.axis_index("pipe"), n_stages)
        local = dp.shard_batch(batch, "data")
        state = pp.init_pp(rng.PRNGKey(0), local.inputs, stage, learning_rate=1e-3)
        if mesh.device.type == "cuda":
            from jax_distributed_tuts_b200.tree import tree_map
            state = state.replace(params=tree_map(lambda t: t.to(mesh.device), state.params), rng=state.rng.to(mesh.device))
            state = state.replace(opt_state=state.tx.init(state.params))
            local = local.to(mesh.device)
        train_step_pp_fn = pp.make_train_step_pp(mesh, stage_model=stage, num_microbatches=MICROBATCHES, pipe_axis="pipe",
                                                 data_axis="data" if n_data > 1 else None)
        metrics = None
        for _ in range(10):
            state, metrics = train_step_pp_fn(state, metrics, local)
        state, final = train_step_pp_fn(state, None, local)
        if rank == 0:
            print(f"mesh data={n_data} x pipe={n_stages}, layers per stage: {[len(r) for r in pp.split_layers(cfg.n_layer, n_stages)]}")
            print_metrics(final, "pp - GPipe")


if __name__ == "__main__":
    launch(worker)

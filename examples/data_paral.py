"""Data-parallel training of the MLP classifier: the counterpart of the reference's data_paral.py.

    python examples/data_paral.py                                   # 8 CPU processes (fake cluster)
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/data_paral.py [--fused]

10 steps on one fixed batch with running metrics, then one more step with fresh metrics whose
loss / accuracy are printed under the title "dp" (data_paral.py:271-277).
"""
import sys

from _common import launch, synthetic_batch

import torch

from jax_distributed_tuts_b200.config import dp_config
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.mlp import Classifier
from jax_distributed_tuts_b200.parallel import dp
from jax_distributed_tuts_b200.state import Batch
from jax_distributed_tuts_b200.utils import get_num_params, print_metrics

FUSED = "--fused" in sys.argv


def worker(rank, world):
    CONFIG = dp_config()
    mesh = Mesh((world,), (CONFIG.data_axis_name,))
    model_dp = Classifier(CONFIG.model)
    model_init_rng, batch = synthetic_batch(CONFIG)
    with mesh:
        state_dp = dp.init_dp(model_init_rng, batch.inputs, model_dp)
        local = dp.shard_batch(batch, CONFIG.data_axis_name)
        if FUSED and torch.cuda.is_available():
            from jax_distributed_tuts_b200.engine import MLPEngine
            eng = MLPEngine(CONFIG.model, mesh, mode="dp", batch_per_rank=local.inputs.shape[0],
                            lr=CONFIG.optimizer.learning_rate, num_minibatches=CONFIG.optimizer.num_minibatches,
                            init_params=state_dp.params)
            state_dp = eng.make_state(model_dp, state_dp.rng)
            local = Batch(local.inputs.pin_memory(), local.labels.pin_memory())
        elif mesh.device.type == "cuda":
            from jax_distributed_tuts_b200.tree import tree_map
            state_dp = state_dp.replace(params=tree_map(lambda t: t.to(mesh.device), state_dp.params),
                                        rng=state_dp.rng.to(mesh.device))
            state_dp = state_dp.replace(opt_state=state_dp.tx.init(state_dp.params))
            local = local.to(mesh.device)
        train_step_dp_fn = dp.make_train_step_dp(mesh, num_minibatches=CONFIG.optimizer.num_minibatches,
                                                 data_axis_name=CONFIG.data_axis_name)
        metrics_dp = None
        for _ in range(10):
            state_dp, metrics_dp = train_step_dp_fn(state_dp, metrics_dp, local)
        state_dp, final_metrics_dp = train_step_dp_fn(state_dp, None, local)
        if rank == 0:
            print(f"devices: {world}, parameters: {get_num_params(state_dp)}")
            print_metrics(final_metrics_dp, title="dp")


if __name__ == "__main__":
    launch(worker)

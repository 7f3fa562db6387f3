"""FSDP-style parameter sharding of the MLP classifier: the counterpart of the reference's
param_sharding.py. 15 steps + 1 with fresh metrics, printed as "FSDP - Final metrics"
(param_sharding.py:389-397). ``--fused`` uses the fused engine on GPUs (all-gather inside the GEMM,
reduce-scatter fused with the sharded AdamW)."""
import sys

from _common import launch, synthetic_batch

import torch

from jax_distributed_tuts_b200.config import fsdp_config
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.mlp import Classifier
from jax_distributed_tuts_b200.parallel import dp, fsdp
from jax_distributed_tuts_b200.state import Batch
from jax_distributed_tuts_b200.tree import get_partition_spec
from jax_distributed_tuts_b200.utils import print_metrics

FUSED = "--fused" in sys.argv


def worker(rank, world):
    config = fsdp_config()          # includes min_weight_size = 2**4 (param_sharding.py:244-246)
    axis = config.model.data_axis_name
    mesh = Mesh((world,), (axis,))
    sharded_dense = lambda m: fsdp.shard_module_params(m, axis_name=axis, min_weight_size=config.model.min_weight_size)
    model_fsdp = Classifier(config.model, dense_wrapper=sharded_dense)
    model_init_rng, batch = synthetic_batch(config)
    with mesh:
        local = dp.shard_batch(batch, axis)
        if FUSED and torch.cuda.is_available():
            from jax_distributed_tuts_b200.engine import MLPEngine
            full = dp.init_dp(model_init_rng, batch.inputs, Classifier(config.model))
            eng = MLPEngine(config.model, mesh, mode="fsdp", batch_per_rank=local.inputs.shape[0], lr=config.model.lr,
                            num_minibatches=config.num_minibatches, init_params=full.params,
                            min_weight_size=config.model.min_weight_size)
            state_fsdp = eng.make_state(model_fsdp, full.rng)
            local = Batch(local.inputs.pin_memory(), local.labels.pin_memory())
        else:
            state_fsdp = fsdp.init_model(model_init_rng, batch.inputs, model_fsdp, learning_rate=config.model.lr)
        state_fsdp_specs = get_partition_spec(state_fsdp)      # param_sharding.py:266
        train_step_fsdp_fn = fsdp.make_train_step_fsdp(mesh, num_minibatches=config.num_minibatches, data_axis_name=axis)
        metrics_fsdp = None
        for _ in range(15):
            state_fsdp, metrics_fsdp = train_step_fsdp_fn(state_fsdp, metrics_fsdp, local)
        state_fsdp, final_metrics_fsdp = train_step_fsdp_fn(state_fsdp, None, local)
        if rank == 0:
            print("partition specs:", {k: dict(v) for k, v in state_fsdp_specs.params.items()})
            print_metrics(final_metrics_fsdp, "FSDP - Final metrics")


if __name__ == "__main__":
    launch(worker)

from .dp import fold_rng_over_axis, init_dp, loss_fn, make_train_step_dp, train_step_dp  # noqa: F401
from .fsdp import (gather_arr_mean_grads, gather_params, init_model, make_train_step_fsdp,  # noqa: F401
                   shard_module_params, shard_params, sync_gradients, train_step_fsdp)

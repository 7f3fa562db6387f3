"""Shared bits of the example drivers: where to run (CPU fake cluster, or the GPUs torchrun gave us)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from jax_distributed_tuts_b200 import rng
from jax_distributed_tuts_b200.launch import init_from_env, sim_multiCPU_dev
from jax_distributed_tuts_b200.state import Batch


def synthetic_batch(cfg, device="cpu"):
    """One fixed synthetic batch, as in the reference (data_paral.py:113-124; labels are the
    intended uniform integers)."""
    key = rng.PRNGKey(cfg.seed)
    model_rng, data_rng, label_rng = rng.split(key, 3)
    x = torch.randn(cfg.data.batch_size, cfg.data.input_size, generator=rng.uniform_generator(data_rng))
    y = torch.randint(0, cfg.data.num_classes, (cfg.data.batch_size,), generator=rng.uniform_generator(label_rng))
    return model_rng, Batch(x.to(device), y.to(device))


def launch(worker, default_devices=8):
    """Under torchrun: run on this rank. Otherwise: the reference's fake cluster, N CPU processes."""
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ:
        rank, world, device = init_from_env()
        worker(rank, world)
    else:
        n = int(os.environ.get("B200DIST_DEVICES", default_devices))
        sim_multiCPU_dev(worker, device_count=n)

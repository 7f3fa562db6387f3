"""Worker functions for the multi-process (gloo) tests; must be importable for mp spawn."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from jax_distributed_tuts_b200 import rng
from jax_distributed_tuts_b200.comm import collectives as C
from jax_distributed_tuts_b200.config import dp_config, fsdp_config
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.mlp import Classifier
from jax_distributed_tuts_b200.parallel import dp, fsdp
from jax_distributed_tuts_b200.state import Batch
from jax_distributed_tuts_b200.tree import get_partition_spec, is_partitioned, tree_leaves, unbox


def make_batch(cfg):
    key = rng.PRNGKey(cfg.seed)
    model_rng, data_rng, label_rng = rng.split(key, 3)
    inputs = torch.randn(cfg.data.batch_size, cfg.data.input_size, generator=rng.uniform_generator(data_rng))
    labels = torch.randint(0, cfg.data.num_classes, (cfg.data.batch_size,),
                           generator=rng.uniform_generator(label_rng))
    return model_rng, Batch(inputs, labels)


def collectives_worker(rank, world):
    mesh = Mesh((world,), ("data",))
    out = {}
    with mesh:
        x = torch.arange(4, dtype=torch.float32) + rank
        out["psum"] = C.psum(x, "data")
        out["pmean"] = C.pmean(x, "data")
        out["idx"] = C.axis_index("data")
        out["size"] = C.axis_size("data")
        sh = torch.full((2, 3), float(rank))
        out["ag0"] = C.all_gather(sh, "data", axis=0)
        out["ag1"] = C.all_gather(sh, "data", axis=1)
        full = torch.arange(world * 2 * 3, dtype=torch.float32).reshape(world * 2, 3) * (rank + 1)
        out["rs0"] = C.psum_scatter(full, "data", scatter_dimension=0)
        full1 = torch.arange(2 * world * 3, dtype=torch.float32).reshape(2, world * 3) * (rank + 1)
        out["rs1"] = C.psum_scatter(full1, "data", scatter_dimension=1)
        out["perm"] = C.ppermute(torch.full((2,), float(rank)), "data", [(i, (i + 1) % world) for i in range(world)])
        out["perm_open"] = C.ppermute(torch.full((2,), float(rank + 1)), "data", [(i, i + 1) for i in range(world - 1)])
        out["scalar"] = C.psum(3, "data")
    return out


def mesh2d_worker(rank, world):
    mesh = Mesh((2, world // 2), ("data", "fsdp"))
    with mesh:
        x = torch.tensor([float(rank)])
        return {"coords": mesh.coords, "sum_data": C.psum(x, "data"), "sum_fsdp": C.psum(x, "fsdp"),
                "sum_both": C.psum(x, ("data", "fsdp")), "idx_both": mesh.axis_index(("data", "fsdp"))}


def dp_worker(rank, world, steps, dropout):
    cfg = dp_config()
    cfg.model.dropout_rate = dropout
    mesh = Mesh((world,), ("data",))
    model = Classifier(cfg.model)
    model_rng, batch = make_batch(cfg)
    with mesh:
        state = dp.init_dp(model_rng, batch.inputs, model)
        step = dp.make_train_step_dp(mesh, num_minibatches=cfg.optimizer.num_minibatches)
        local = dp.shard_batch(batch)
        metrics = None
        for _ in range(steps):
            state, metrics = step(state, metrics, local)
    return {"params": state.params, "metrics": metrics, "step": state.step}


def fsdp_worker(rank, world, steps, lr, min_weight_size=16):
    cfg = fsdp_config()
    cfg.model.dropout_rate = 0.0
    cfg.model.min_weight_size = min_weight_size
    mesh = Mesh((world,), ("data",))
    wrap = lambda m: fsdp.shard_module_params(m, axis_name="data", min_weight_size=cfg.model.min_weight_size)
    model = Classifier(cfg.model, dense_wrapper=wrap)
    model_rng, batch = make_batch(cfg)
    with mesh:
        state = fsdp.init_model(model_rng, batch.inputs, model, learning_rate=lr)
        specs = get_partition_spec(state)
        step = fsdp.make_train_step_fsdp(mesh, num_minibatches=cfg.num_minibatches)
        local = dp.shard_batch(batch)
        metrics = None
        for _ in range(steps):
            state, metrics = step(state, metrics, local)
        full = fsdp.unshard_params(state.params)
        local_shapes = {k: {kk: tuple((vv.value if is_partitioned(vv) else vv).shape) for kk, vv in v.items()}
                        for k, v in state.params.items()}
        names = {k: {kk: (vv.names if is_partitioned(vv) else None) for kk, vv in v.items()}
                 for k, v in state.params.items()}
        mu_names = {k: {kk: (vv.names if is_partitioned(vv) else None) for kk, vv in v.items()}
                    for k, v in state.opt_state[0].mu.items()}
    return {"params": full, "metrics": metrics, "local_shapes": local_shapes, "names": names,
            "mu_names": mu_names, "spec_params": specs.params}


def dp_lr_worker(rank, world, steps, lr):
    """DP with the FSDP config (same seed/lr) for the FSDP == DP equivalence test."""
    from jax_distributed_tuts_b200.optim import adamw
    cfg = fsdp_config()
    cfg.model.dropout_rate = 0.0
    mesh = Mesh((world,), ("data",))
    model = Classifier(cfg.model)
    model_rng, batch = make_batch(cfg)
    with mesh:
        state = dp.init_dp(model_rng, batch.inputs, model, optimizer=adamw(lr))
        step = dp.make_train_step_dp(mesh, num_minibatches=cfg.num_minibatches)
        local = dp.shard_batch(batch)
        metrics = None
        for _ in range(steps):
            state, metrics = step(state, metrics, local)
    return {"params": state.params, "metrics": metrics}


def pp_worker(rank, world, n_stages, steps, microbatches):
    """GPipe over `n_stages` ranks (x `world // n_stages` data replicas) on a 4-layer tiny GPT-2."""
    from jax_distributed_tuts_b200.models.gpt2 import GPT2, gpt2_config
    from jax_distributed_tuts_b200.parallel import pp
    cfg = gpt2_config("tiny", n_layer=4)
    n_data = world // n_stages
    mesh = Mesh((n_data, n_stages), ("data", "pipe"))
    model = GPT2(cfg)
    g = torch.Generator().manual_seed(0)
    tok = torch.randint(0, cfg.vocab_size, (4 * n_data, 129), generator=g)
    batch = Batch(tok[:, :128].contiguous(), tok[:, 1:].contiguous())
    with mesh:
        stage = pp.GPT2Stage(model, mesh.axis_index("pipe"), n_stages)
        local = dp.shard_batch(batch, "data")
        state = pp.init_pp(rng.PRNGKey(7), local.inputs, stage, learning_rate=1e-3)
        step = pp.make_train_step_pp(mesh, stage_model=stage, num_microbatches=microbatches, pipe_axis="pipe",
                                     data_axis="data" if n_data > 1 else None)
        metrics = None
        for _ in range(steps):
            state, metrics = step(state, metrics, local)
    return {"params": state.params, "metrics": metrics, "stage": mesh.axis_index("pipe"),
            "data": mesh.axis_index("data")}


def fsdp_ckpt_worker(rank, world, path):
    from jax_distributed_tuts_b200 import checkpoint as ckpt
    cfg = fsdp_config()
    cfg.model.dropout_rate = 0.0
    mesh = Mesh((world,), ("data",))
    wrap = lambda m: fsdp.shard_module_params(m, axis_name="data", min_weight_size=16)
    model = Classifier(cfg.model, dense_wrapper=wrap)
    model_rng, batch = make_batch(cfg)
    with mesh:
        state = fsdp.init_model(model_rng, batch.inputs, model, learning_rate=1e-3)
        step = fsdp.make_train_step_fsdp(mesh, num_minibatches=2)
        state, _ = step(state, None, dp.shard_batch(batch))
        ckpt.save_checkpoint(path, state, mesh=mesh)
        import torch.distributed as dist
        dist.barrier()
        loaded = ckpt.load_checkpoint(path)
        restored = ckpt.restore_params(state.params, loaded["params"], mesh=mesh)
        k0, k1 = state.params["input_dense"]["kernel"], restored["input_dense"]["kernel"]
        same = torch.equal(k0.value, k1.value) and k1.names == k0.names
        same &= torch.equal(state.params["output_dense"]["bias"], restored["output_dense"]["bias"])
    return {"reload_equal": bool(same), "local_shape": tuple(k1.value.shape)}


def fsdp2d_worker(rank, world, steps, lr):
    """2-D mesh (data=2, fsdp=world/2): parameters sharded over `fsdp`, gradients additionally
    averaged over `data` by sync_gradients (param_sharding.py:293-322)."""
    cfg = fsdp_config()
    cfg.model.dropout_rate = 0.0
    mesh = Mesh((2, world // 2), ("data", "fsdp"))
    wrap = lambda m: fsdp.shard_module_params(m, axis_name="fsdp", min_weight_size=16)
    model = Classifier(cfg.model, dense_wrapper=wrap)
    model_rng, batch = make_batch(cfg)
    with mesh:
        state = fsdp.init_model(model_rng, batch.inputs, model, learning_rate=lr)
        # every rank of the (data, fsdp) grid takes a distinct slice of the batch
        n, i = mesh.size, mesh.axis_index(("data", "fsdp"))
        per = batch.inputs.shape[0] // n
        local = batch.slice(i * per, (i + 1) * per)
        loss_fn = dp.make_loss_fn("fsdp")
        step = fsdp.make_train_step_fsdp(mesh, num_minibatches=cfg.num_minibatches, data_axis_name="fsdp",
                                         loss_fn=loss_fn, sync_axes=("data", "fsdp"))
        metrics = None
        for _ in range(steps):
            state, metrics = step(state, metrics, local)
        full = fsdp.unshard_params(state.params)
        names = {k: {kk: (vv.names if is_partitioned(vv) else None) for kk, vv in v.items()}
                 for k, v in state.params.items()}
    return {"params": full, "metrics": metrics, "names": names, "coords": mesh.coords}


def pp_ckpt_worker(rank, world, path):
    """2-stage GPipe, one step, global checkpoint merged over the pipe axis, restored into each stage."""
    from jax_distributed_tuts_b200 import checkpoint as ckpt
    from jax_distributed_tuts_b200.models.gpt2 import GPT2, gpt2_config
    from jax_distributed_tuts_b200.parallel import pp
    import torch.distributed as dist
    cfg = gpt2_config("tiny", n_layer=4)
    mesh = Mesh((1, world), ("data", "pipe"))
    model = GPT2(cfg)
    g = torch.Generator().manual_seed(0)
    tok = torch.randint(0, cfg.vocab_size, (4, 129), generator=g)
    batch = Batch(tok[:, :128].contiguous(), tok[:, 1:].contiguous())
    with mesh:
        stage = pp.GPT2Stage(model, mesh.axis_index("pipe"), world)
        state = pp.init_pp(rng.PRNGKey(7), batch.inputs, stage, learning_rate=1e-3)
        step = pp.make_train_step_pp(mesh, stage_model=stage, num_microbatches=2, pipe_axis="pipe")
        state, _ = step(state, None, batch)
        refused = False
        try:
            ckpt.save_checkpoint(path, state, mesh=mesh)
        except ValueError:
            refused = True
        ckpt.save_checkpoint(path, state, mesh=mesh, pipe_axis="pipe")
        dist.barrier()
        loaded = ckpt.load_checkpoint(path)
        restored = ckpt.restore_params(state.params, loaded["params"], mesh=mesh)
        same = all(torch.equal(a, b) for a, b in zip(tree_leaves(state.params), tree_leaves(restored)))
    return {"refused_without_axis": refused, "reload_equal": bool(same), "top_keys": sorted(loaded["params"].keys()),
            "mu_keys": sorted(loaded["opt_state"]["0"]["mu"].keys())}


def fused_backend_fallback_worker(rank, world):
    """set_backend("fused") on CPU ranks: the fused kernels need GPUs, the calls fall back to the axis' gloo group."""
    mesh = Mesh((world,), ("data",))
    C.set_backend("fused")
    try:
        with mesh:
            t = torch.full((5,), float(rank + 1))
            s = C.psum(t, "data")
            g = C.all_gather(torch.tensor([rank], dtype=torch.int64), "data", axis=0)
            rs = C.psum_scatter(torch.arange(world * 2, dtype=torch.float32) + rank, "data", scatter_dimension=0)
    finally:
        C.set_backend("auto")
    return {"psum": s.tolist(), "gather": g.tolist(), "scatter": rs.tolist(), "fused_calls": C.fused_calls}


def dying_peer_worker(rank, world):
    """Rank 1 dies before the collective; the others must get an error out of it, not wait forever."""
    import torch.distributed as dist
    if rank == 1:
        os._exit(17)
    t = torch.ones(4)
    dist.all_reduce(t)          # raises once the time-out set by the launcher expires (or the connection drops)
    return {"sum": t.tolist()}

"""CPU plumbing tier: N gloo processes stand in for N devices (the role of the reference's
``sim_multiCPU_dev`` fake cluster, util.py:31-38). DP == single process, FSDP == DP."""
import pytest
import torch

from jax_distributed_tuts_b200.launch import sim_multiCPU_dev
from jax_distributed_tuts_b200.tree import P, tree_leaves

import _workers as W


def _close(a, b, lr=1e-3, steps=3):
    d = (a - b).abs()
    return d.max() <= 2.2 * lr * steps and d.mean() <= 0.05 * lr * steps


def test_collectives_by_axis_name():
    world = 4
    res = sim_multiCPU_dev(W.collectives_worker, world)
    base = torch.arange(4, dtype=torch.float32)
    for r, out in enumerate(res):
        assert torch.equal(out["psum"], base * world + sum(range(world)))
        assert torch.allclose(out["pmean"], base + sum(range(world)) / world)
        assert out["idx"] == r and out["size"] == world and out["scalar"] == 3 * world
        assert out["ag0"].shape == (2 * world, 3) and torch.equal(out["ag0"][2 * r], torch.full((3,), float(r)))
        assert out["ag1"].shape == (2, 3 * world) and torch.equal(out["ag1"][:, 3 * r], torch.full((2,), float(r)))
        scale = sum(i + 1 for i in range(world))
        full = torch.arange(world * 2 * 3, dtype=torch.float32).reshape(world * 2, 3) * scale
        assert torch.equal(out["rs0"], full[2 * r:2 * r + 2])
        full1 = torch.arange(2 * world * 3, dtype=torch.float32).reshape(2, world * 3) * scale
        assert torch.equal(out["rs1"], full1[:, 3 * r:3 * r + 3])
        assert torch.equal(out["perm"], torch.full((2,), float((r - 1) % world)))
        assert torch.equal(out["perm_open"], torch.full((2,), float(r) if r > 0 else 0.0))


def test_mesh_2d_axis_groups():
    res = sim_multiCPU_dev(W.mesh2d_worker, 4)
    # ranks laid out row-major over (data=2, fsdp=2)
    assert [r["coords"] for r in res] == [(0, 0), (0, 1), (1, 0), (1, 1)]
    assert [r["sum_data"].item() for r in res] == [2.0, 4.0, 2.0, 4.0]
    assert [r["sum_fsdp"].item() for r in res] == [1.0, 1.0, 5.0, 5.0]
    assert all(r["sum_both"].item() == 6.0 for r in res)
    assert [r["idx_both"] for r in res] == [0, 1, 2, 3]


def test_dp_matches_single_process():
    steps = 3
    multi = sim_multiCPU_dev(W.dp_worker, 4, args=(steps, 0.0))
    single = sim_multiCPU_dev(W.dp_worker, 1, args=(steps, 0.0))[0]
    for r in multi[1:]:   # replicas stay identical
        for a, b in zip(tree_leaves(multi[0]["params"]), tree_leaves(r["params"])):
            assert torch.equal(a, b)
    # Adam turns bf16 rounding noise on near-zero gradients into +-lr steps, so compare in the mean
    for a, b in zip(tree_leaves(multi[0]["params"]), tree_leaves(single["params"])):
        assert _close(a, b), (a - b).abs().mean()
    assert multi[0]["metrics"]["loss"][1] == 128 * steps
    assert abs(multi[0]["metrics"]["loss"][0].item() - single["metrics"]["loss"][0].item()) < 0.05 * steps
    assert multi[0]["step"] == steps


def test_dp_dropout_streams_differ_per_device():
    a = sim_multiCPU_dev(W.dp_worker, 2, args=(1, 0.1))
    b = sim_multiCPU_dev(W.dp_worker, 2, args=(1, 0.0))
    d = (a[0]["params"]["input_dense"]["kernel"] - b[0]["params"]["input_dense"]["kernel"]).abs().max()
    assert d > 0  # dropout changes the update; replicas still agree
    for x, y in zip(tree_leaves(a[0]["params"]), tree_leaves(a[1]["params"])):
        assert torch.equal(x, y)


def test_fsdp_shards_like_reference_and_matches_dp():
    world, steps, lr = 8, 2, 1e-3
    res = sim_multiCPU_dev(W.fsdp_worker, world, args=(steps, lr))
    r0 = res[0]
    # shard decisions with min_weight_size = 2**4 on 8 devices
    assert r0["local_shapes"]["input_dense"]["kernel"] == (98, 512)
    assert r0["local_shapes"]["input_dense"]["bias"] == (64,)
    assert r0["local_shapes"]["output_dense"]["kernel"] == (64, 10)
    assert r0["local_shapes"]["output_dense"]["bias"] == (10,)
    assert r0["names"]["input_dense"]["kernel"] == ("data", None) and r0["names"]["output_dense"]["bias"] is None
    assert r0["mu_names"]["input_dense"]["kernel"] == ("data", None)      # sharded optimizer state
    assert r0["spec_params"]["input_dense"]["kernel"] == P("data", None)
    assert r0["spec_params"]["output_dense"]["bias"] == P()
    dp_res = sim_multiCPU_dev(W.dp_lr_worker, world, args=(steps, lr))[0]
    for a, b in zip(tree_leaves(r0["params"]), tree_leaves(dp_res["params"])):
        assert a.shape == b.shape
        assert _close(a, b), (a - b).abs().mean()
    assert abs(r0["metrics"]["loss"][0].item() - dp_res["metrics"]["loss"][0].item()) < 1e-2


def test_fsdp_default_min_weight_size_only_shards_big_kernel():
    res = sim_multiCPU_dev(W.fsdp_worker, 2, args=(1, 1e-4, 2 ** 18))
    names = res[0]["names"]
    assert names["input_dense"]["kernel"] == ("data", None)
    assert names["input_dense"]["bias"] is None and names["output_dense"]["kernel"] is None


def test_fsdp_on_2d_mesh_matches_dp():
    """BASELINE.json config 5 (2-D mesh data x param sharding), generic tier on the CPU fake cluster."""
    world, steps, lr = 8, 2, 1e-3
    res = sim_multiCPU_dev(W.fsdp2d_worker, world, args=(steps, lr))
    r0 = res[0]
    assert r0["names"]["input_dense"]["kernel"] == ("fsdp", None)      # 784 / 4 = 196 rows per shard
    dp_res = sim_multiCPU_dev(W.dp_lr_worker, world, args=(steps, lr))[0]
    for a, b in zip(tree_leaves(r0["params"]), tree_leaves(dp_res["params"])):
        assert _close(a, b), (a - b).abs().mean()
    for r in res[1:]:       # all 8 ranks reassemble the same parameters
        for a, b in zip(tree_leaves(r0["params"]), tree_leaves(r["params"])):
            assert torch.equal(a, b)
    assert r0["metrics"]["loss"][1] == 128 * steps


def test_fused_collective_backend_falls_back_on_cpu():
    res = sim_multiCPU_dev(W.fused_backend_fallback_worker, 2)
    for r, out in enumerate(res):
        assert out["psum"] == [3.0] * 5 and out["gather"] == [0, 1] and out["fused_calls"] == 0
        assert out["scatter"] == [float(2 * (2 * r) + 1), float(2 * (2 * r + 1) + 1)]


def test_a_dead_rank_fails_the_job_instead_of_hanging_it():
    import time
    t0 = time.time()
    with pytest.raises(RuntimeError):
        sim_multiCPU_dev(W.dying_peer_worker, 3, timeout_s=8.0)
    assert time.time() - t0 < 90.0

"""Checkpoint layout and round trips: the reference-shaped state dict, Flax msgpack wire format,
safetensors, sharded save -> unsharded file -> sharded load, and shard/gather properties."""
import os

import numpy as np
import pytest
import torch
from hypothesis import given, settings
from hypothesis import strategies as st

from jax_distributed_tuts_b200 import checkpoint as ckpt
from jax_distributed_tuts_b200 import rng
from jax_distributed_tuts_b200.config import dp_config
from jax_distributed_tuts_b200.launch import sim_multiCPU_dev
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.mlp import Classifier
from jax_distributed_tuts_b200.parallel import dp
from jax_distributed_tuts_b200.parallel.fsdp import choose_shard_dim
from jax_distributed_tuts_b200.state import Batch

import _workers as W


def _dp_state():
    cfg = dp_config()
    mesh = Mesh((1,), ("data",))
    model = Classifier(cfg.model)
    x = torch.randn(8, 784)
    with mesh:
        state = dp.init_dp(rng.PRNGKey(0), x, model)
        step = dp.make_train_step_dp(mesh, num_minibatches=2)
        state, _ = step(state, None, Batch(x, torch.randint(0, 10, (8,))))
    return state


@pytest.mark.parametrize("ext", [".msgpack", ".safetensors"])
def test_state_dict_layout_and_round_trip(tmp_path, ext):
    state = _dp_state()
    sd = ckpt.state_dict(state)
    assert set(sd) == {"step", "params", "opt_state", "rng"}
    assert set(sd["params"]) == {"input_dense", "output_dense"}
    assert sd["params"]["input_dense"]["kernel"].shape == (784, 512) and sd["params"]["input_dense"]["kernel"].dtype == np.float32
    assert sd["params"]["output_dense"]["kernel"].shape == (512, 10)
    assert set(sd["opt_state"]) == {"0", "1", "2"} and set(sd["opt_state"]["0"]) == {"count", "mu", "nu"}
    assert sd["opt_state"]["1"] == {} and int(sd["opt_state"]["0"]["count"]) == 1 and int(sd["step"]) == 1
    path = str(tmp_path / ("ck" + ext))
    ckpt.save_checkpoint(path, state, rank=0)
    loaded = ckpt.load_checkpoint(path)
    for a in ("input_dense", "output_dense"):
        for b in ("kernel", "bias"):
            assert np.array_equal(loaded["params"][a][b], sd["params"][a][b])
            assert np.array_equal(loaded["opt_state"]["0"]["mu"][a][b], sd["opt_state"]["0"]["mu"][a][b])
    restored = ckpt.restore_params(state.params, loaded["params"])
    assert torch.equal(restored["input_dense"]["kernel"], state.params["input_dense"]["kernel"])


def test_msgpack_is_flax_wire_format():
    import msgpack
    arr = np.arange(6, dtype=np.float32).reshape(2, 3)
    blob = ckpt.msgpack_serialize({"params": {"w": arr}, "step": np.asarray(3, np.int32)})
    raw = msgpack.unpackb(blob, raw=False)                       # without our hook: ExtType(1, ...)
    ext = raw["params"]["w"]
    assert isinstance(ext, msgpack.ExtType) and ext.code == 1
    shape, dtype_name, buf = msgpack.unpackb(ext.data, raw=False)
    assert shape == [2, 3] and dtype_name == "float32" and np.array_equal(np.frombuffer(buf, np.float32).reshape(2, 3), arr)
    back = ckpt.msgpack_restore(blob)
    assert np.array_equal(back["params"]["w"], arr) and int(back["step"]) == 3
    # a Partitioned leaf as Flax would serialise it ({"value": array}) is unwrapped on load
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "flax.msgpack")
        open(p, "wb").write(ckpt.msgpack_serialize({"params": {"w": {"value": arr}}}))
        assert np.array_equal(ckpt.load_checkpoint(p)["params"]["w"], arr)


def test_fsdp_checkpoint_is_global_and_reloads_sharded(tmp_path):
    path = str(tmp_path / "fsdp.msgpack")
    res = sim_multiCPU_dev(W.fsdp_ckpt_worker, 4, args=(path,))
    loaded = ckpt.load_checkpoint(path)
    assert loaded["params"]["input_dense"]["kernel"].shape == (784, 512)      # full shapes in the file
    assert loaded["opt_state"]["0"]["mu"]["input_dense"]["kernel"].shape == (784, 512)
    for r in res:
        assert r["reload_equal"] and r["local_shape"] == (196, 512)
    # the same file loads into an unsharded (DP) template
    state = _dp_state()
    restored = ckpt.restore_params(state.params, loaded["params"])
    assert np.array_equal(restored["input_dense"]["kernel"].numpy(), loaded["params"]["input_dense"]["kernel"])


@settings(max_examples=60, deadline=None)
@given(shape=st.lists(st.integers(1, 48), min_size=1, max_size=3), n=st.sampled_from([1, 2, 4, 8]))
def test_shard_then_gather_is_identity(shape, n):
    """gather(shard(x)) == x for the reference's shard rule on random shapes / axis sizes."""
    x = torch.arange(int(np.prod(shape)), dtype=torch.float32).reshape(shape)
    dim = choose_shard_dim(tuple(shape), (None,) * len(shape), n)
    if dim is None:
        assert all(s % n != 0 for s in shape)
        return
    # largest divisible dim, ties resolved like np.argsort(shape)[::-1]
    order = list(np.argsort(shape)[::-1])
    assert dim == next(i for i in order if shape[i] % n == 0)
    per = shape[dim] // n
    shards = [x.narrow(dim, i * per, per) for i in range(n)]
    assert torch.equal(torch.cat(shards, dim=dim), x)

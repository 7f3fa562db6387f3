"""Checkpointing in the layout the reference's state would serialise to.

The reference never saves anything, but its state pytree fixes the layout a Flax user would get
from ``flax.serialization``: ``{"step", "params", "opt_state": {"0": {"count", "mu", "nu"}, "1": {},
"2": {}}, "rng"}`` with parameters ``{"input_dense": {"kernel" (784, 512), "bias"}, "output_dense":
{...}}`` in fp32, (in, out) kernels (data_paral.py:86-98,137-143). Checkpoints here always hold the
*global, unsharded* arrays — the same tensors for DP, FSDP, 2-D mesh and pipeline runs — so a file
written by one strategy loads into any other; sharded runs gather on save and slice on load with
the same shard rule as ``shard_params`` (param_sharding.py:89-113); pipeline runs merge the stages'
trees over the ``pipe`` axis on save (``pipe_axis=``; the tied ``wte`` is stored once) and each stage
picks its own blocks on load (``restore_params`` with the stage's tree as template).

Two containers:
  * ``.msgpack``     Flax's wire format: msgpack map with ndarrays as ExtType 1 holding
                     ``packb((shape, dtype_name, raw_bytes))`` — readable by ``flax.serialization.
                     msgpack_restore`` and vice versa.
  * ``.safetensors`` flat ``a/b/c`` keys, plus the per-leaf partition names in the metadata.
"""
from __future__ import annotations

import json
import os
from typing import Any, Dict, Optional, Tuple

import msgpack
import numpy as np
import torch

from .mesh import Mesh
from .rng import Key
from .tree import Pytree, is_partitioned, tree_leaves_with_path, tree_map

_EXT_NDARRAY = 1
_EXT_NPSCALAR = 3


# ------------------------------------------------------------------------------------------------
# state <-> nested dict of numpy arrays
# ------------------------------------------------------------------------------------------------
def _to_numpy(x) -> np.ndarray:
    if isinstance(x, torch.Tensor):
        t = x.detach().cpu()
        if t.dtype == torch.bfloat16:
            t = t.float()
        return t.contiguous().numpy()
    return np.asarray(x)


def _plain(tree: Pytree) -> Any:
    """Nested dict / list -> nested dict with string keys (Flax ``to_state_dict`` conventions:
    tuples and lists become ``{"0": ..., "1": ...}``, dataclasses their field names)."""
    import dataclasses
    if is_partitioned(tree):
        return _to_numpy(tree.value)
    if isinstance(tree, Key):
        return _to_numpy(tree.base)
    if isinstance(tree, dict):
        return {str(k): _plain(v) for k, v in tree.items()}
    if isinstance(tree, (list, tuple)):
        return {str(i): _plain(v) for i, v in enumerate(tree)}
    if dataclasses.is_dataclass(tree) and not isinstance(tree, type):
        return {f.name: _plain(getattr(tree, f.name)) for f in dataclasses.fields(tree)}
    if tree is None:
        return {}
    return _to_numpy(tree)


def partition_names(params: Pytree) -> Dict[str, Tuple[Optional[str], ...]]:
    return {p: tuple(v.names) for p, v in tree_leaves_with_path(params, is_leaf=is_partitioned) if is_partitioned(v)}


def _merge_trees(trees) -> Dict[str, Any]:
    """Union of nested dicts by key; where several stages hold the same leaf (the tied ``wte`` lives
    on the first and on the last pipeline stage) the first one is kept."""
    out: Dict[str, Any] = {}
    for t in trees:
        for k, v in t.items():
            if isinstance(v, dict):
                out[k] = _merge_trees([out.get(k, {}), v]) if isinstance(out.get(k), dict) else _merge_trees([v])
            else:
                out.setdefault(k, v)
    return out


def merge_over_pipe(sd: Dict[str, Any], mesh: Mesh, pipe_axis: str) -> Dict[str, Any]:
    """A pipeline stage's state dict holds only the blocks that stage owns. Gathers the stages' dicts over
    the ``pipe`` axis (every rank of the axis calls this) and returns the union: the whole model's
    parameters and Adam moments, as a single-device run would have saved them."""
    import torch.distributed as dist
    n = mesh.axis_size(pipe_axis)
    if n == 1:
        return sd
    parts = [None] * n
    dist.all_gather_object(parts, {"params": sd["params"], "opt": sd["opt_state"]["0"]}, group=mesh.group(pipe_axis))
    out = dict(sd)
    out["params"] = _merge_trees([p["params"] for p in parts])
    adam = dict(sd["opt_state"]["0"])
    adam["mu"] = _merge_trees([p["opt"]["mu"] for p in parts])
    adam["nu"] = _merge_trees([p["opt"]["nu"] for p in parts])
    out["opt_state"] = dict(sd["opt_state"], **{"0": adam})
    return out


def state_dict(state, *, mesh: Optional[Mesh] = None, pipe_axis: Optional[str] = None) -> Dict[str, Any]:
    """Global (unsharded) state dict of a TrainState. Sharded leaves are gathered over their named
    axes; states owned by a fused engine read the engine's buffers; a pipeline stage's state
    (``pipe_axis`` names the mesh axis the stages sit on) is merged with the other stages' into the
    whole model."""
    if pipe_axis is not None and mesh is not None and pipe_axis in mesh.axis_names:
        return merge_over_pipe(state_dict(state, mesh=mesh), mesh, pipe_axis)
    from .parallel.fsdp import unshard_params
    eng = getattr(state, "engine", None)
    if eng is not None:
        # same container as the generic tiers: mu / nu shaped like the parameters, global layout
        params = eng.params_tree(gather=True)
        og = eng.opt_state_global()
        opt = {"0": {"count": _to_numpy(og["count"]).reshape(()), "mu": _plain(og["mu"]), "nu": _plain(og["nu"])},
               "1": {}, "2": {}}
        rng_arr = _to_numpy(eng.rng_base)
    else:
        needs_mesh = any(is_partitioned(v) for _, v in tree_leaves_with_path(state.params, is_leaf=is_partitioned))
        params = unshard_params(state.params, mesh) if needs_mesh else state.params
        opt_tree = state.opt_state
        if needs_mesh:
            opt_tree = tree_map(lambda x: x, opt_tree)  # copy structure
            adam = opt_tree[0]
            import dataclasses
            adam = dataclasses.replace(adam, mu=unshard_params(adam.mu, mesh), nu=unshard_params(adam.nu, mesh))
            opt_tree = (adam,) + tuple(opt_tree[1:])
        opt = _plain(opt_tree)
        rng_arr = _to_numpy(state.rng.base) if state.rng is not None else np.zeros(1, np.int64)
    return {"step": np.asarray(int(state.step), dtype=np.int32), "params": _plain(params), "opt_state": opt,
            "rng": rng_arr}


# ------------------------------------------------------------------------------------------------
# Flax msgpack wire format
# ------------------------------------------------------------------------------------------------
def _ext_pack(x):
    if isinstance(x, np.ndarray):
        return msgpack.ExtType(_EXT_NDARRAY, msgpack.packb((list(x.shape), x.dtype.name, x.tobytes("C")), use_bin_type=True))
    if isinstance(x, np.generic):
        return msgpack.ExtType(_EXT_NPSCALAR, msgpack.packb((list(np.asarray(x).shape), np.asarray(x).dtype.name,
                                                              np.asarray(x).tobytes("C")), use_bin_type=True))
    raise TypeError(f"cannot serialise {type(x)}")


def _ext_unpack(code, data):
    if code in (_EXT_NDARRAY, _EXT_NPSCALAR):
        shape, dtype_name, buf = msgpack.unpackb(data, raw=False)
        arr = np.frombuffer(buf, dtype=np.dtype(dtype_name)).reshape(shape)
        return arr.copy() if code == _EXT_NDARRAY else arr.reshape(())[()]
    return msgpack.ExtType(code, data)


def msgpack_serialize(tree: Dict[str, Any]) -> bytes:
    return msgpack.packb(tree, default=_ext_pack, strict_types=True, use_bin_type=True)


def msgpack_restore(data: bytes) -> Dict[str, Any]:
    return msgpack.unpackb(data, ext_hook=_ext_unpack, raw=False)


# ------------------------------------------------------------------------------------------------
# save / load
# ------------------------------------------------------------------------------------------------
def _flatten(tree: Dict[str, Any], prefix: str = "") -> Dict[str, np.ndarray]:
    out: Dict[str, np.ndarray] = {}
    for k, v in tree.items():
        p = f"{prefix}/{k}" if prefix else k
        if isinstance(v, dict):
            out.update(_flatten(v, p))
        else:
            out[p] = np.asarray(v)
    return out


def _unflatten(flat: Dict[str, Any]) -> Dict[str, Any]:
    out: Dict[str, Any] = {}
    for path, v in flat.items():
        node = out
        *parents, last = path.split("/")
        for p in parents:
            node = node.setdefault(p, {})
        node[last] = v
    return out


def save_checkpoint(path: str, state, *, mesh: Optional[Mesh] = None, rank: Optional[int] = None,
                    pipe_axis: Optional[str] = None) -> Optional[str]:
    """Writes the global state. Every rank must call it (sharded leaves are gathered collectively, pipeline
    stages are merged over ``pipe_axis``); only rank 0 touches the file system. The container follows the
    file extension. A state that holds only part of the model (a pipeline stage) is refused unless
    ``pipe_axis`` says how to complete it: a file with one stage's blocks would restore into nothing."""
    if pipe_axis is None and mesh is not None and "pipe" in mesh.axis_names and mesh.axis_size("pipe") > 1 \
            and getattr(state, "engine", None) is None:
        raise ValueError("this mesh has a 'pipe' axis: pass pipe_axis='pipe' so the stages' parameters are merged "
                         "into one global checkpoint (each stage's TrainState holds only its own blocks)")
    sd = state_dict(state, mesh=mesh, pipe_axis=pipe_axis)
    if rank is None:
        import torch.distributed as dist
        rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
    if rank != 0:
        return None
    names = partition_names(state.params) if getattr(state, "engine", None) is None else {}
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    tmp = path + ".tmp"
    if path.endswith(".safetensors"):
        from safetensors.numpy import save_file
        flat = {k: np.ascontiguousarray(v) for k, v in _flatten(sd).items()}
        save_file(flat, tmp, metadata={"format": "b200dist-global-v1",
                                       "partition_names": json.dumps({k: list(v) for k, v in names.items()})})
    else:
        with open(tmp, "wb") as f:
            f.write(msgpack_serialize(sd))
        with open(path + ".names.json", "w") as f:
            json.dump({k: list(v) for k, v in names.items()}, f)
    os.replace(tmp, path)
    return path


def load_checkpoint(path: str) -> Dict[str, Any]:
    """Nested dict of numpy arrays (global layout). Accepts files written by Flax too: a
    ``Partitioned`` leaf serialised as ``{"value": array}`` is unwrapped."""
    if path.endswith(".safetensors"):
        from safetensors.numpy import load_file
        tree = _unflatten(load_file(path))
    else:
        with open(path, "rb") as f:
            tree = msgpack_restore(f.read())

    def unwrap(t):
        if isinstance(t, dict):
            if set(t.keys()) == {"value"} and not isinstance(t["value"], dict):
                return t["value"]
            return {k: unwrap(v) for k, v in t.items()}
        return t

    return unwrap(tree)


def restore_params(template: Pytree, loaded: Dict[str, Any], *, mesh: Optional[Mesh] = None) -> Pytree:
    """Puts loaded global arrays into the structure (and sharding) of ``template``: plain leaves are
    replaced, ``Partitioned`` leaves receive this rank's slice along their named dimensions."""
    flat = _flatten(loaded)

    def put(path, leaf):
        arr = flat[path]
        if is_partitioned(leaf):
            t = torch.from_numpy(np.ascontiguousarray(arr))
            for d, name in enumerate(leaf.names):
                if name is not None:
                    n, i = mesh.axis_size(name), mesh.axis_index(name)
                    per = t.shape[d] // n
                    t = t.narrow(d, i * per, per)
            return leaf.replace(value=t.contiguous().to(leaf.value.device, leaf.value.dtype))
        return torch.from_numpy(np.ascontiguousarray(arr)).to(leaf.device, leaf.dtype)

    paths = iter([p for p, _ in tree_leaves_with_path(template, is_leaf=is_partitioned)])
    return tree_map(lambda leaf: put(next(paths), leaf), template, is_leaf=is_partitioned)


def restore_engine(engine, loaded: Dict[str, Any]) -> None:
    """Puts a loaded global checkpoint (``load_checkpoint``) into a fused engine: parameters,
    Adam moments and step count (each rank keeps its slice / shard), and the RNG key. The file may
    come from a generic-tier TrainState or from an engine of another mode or world size."""
    import torch
    engine.load_params(loaded["params"])
    adam = loaded["opt_state"]["0"]
    engine.load_opt_state_global(adam["mu"], adam["nu"], int(np.asarray(adam["count"]).reshape(-1)[0]))
    engine.rng_base.copy_(torch.from_numpy(np.asarray(loaded["rng"]).astype(np.int64).reshape(-1)[:1]).to(engine.rng_base.device))
    engine.steps_done = int(np.asarray(loaded["step"]).reshape(-1)[0])

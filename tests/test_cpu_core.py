"""CPU tier: pytree / config / RNG / optimizer / accumulation against plain PyTorch."""
import math

import pytest
import torch

import jax_distributed_tuts_b200 as jdt
from jax_distributed_tuts_b200 import rng
from jax_distributed_tuts_b200.accum import accum_grads
from jax_distributed_tuts_b200.config import ConfigDict, apply_overrides, dp_config, fsdp_config
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.mlp import Classifier
from jax_distributed_tuts_b200.optim import adamw
from jax_distributed_tuts_b200.parallel import dp
from jax_distributed_tuts_b200.state import Batch, TrainState
from jax_distributed_tuts_b200.tree import P, Partitioned, get_partition_spec, tree_leaves, tree_leaves_with_path, tree_map
from jax_distributed_tuts_b200.utils.metrics import format_metrics, get_num_params


def test_config_defaults_match_reference():
    c = dp_config()
    assert c.data.batch_size == 128 and c.data.num_classes == 10 and c.data.input_size == 784
    assert c.model.hidden_size == 512 and c.model.dropout_rate == 0.1 and c.model.dtype == torch.bfloat16
    assert c.optimizer.learning_rate == 1e-3 and c.optimizer.num_minibatches == 4 and c.seed == 69
    f = fsdp_config()
    assert f.model.lr == 1e-4 and f.seed == 6969 and f.num_minibatches == 4 and f.model.min_weight_size == 16
    apply_overrides(c, ["--model.hidden_size=1024", "optimizer.learning_rate=0.01"])
    assert c.model.hidden_size == 1024 and c.optimizer.learning_rate == 0.01
    with pytest.raises(KeyError):
        apply_overrides(c, ["model.nope=1"])
    with pytest.raises(AttributeError):
        _ = c.OPTIMIZER_CONFIG  # the reference's typo is an error here too


def test_tree_map_and_partition_spec():
    t = {"a": torch.ones(2), "b": {"c": Partitioned(torch.zeros(4, 2), ("data", None)), "d": None}}
    doubled = tree_map(lambda x: x * 2 if isinstance(x, torch.Tensor) else x.replace(value=x.value + 1), t,
                       is_leaf=lambda x: isinstance(x, Partitioned))
    assert doubled["a"].sum() == 4 and doubled["b"]["c"].value.sum() == 8 and doubled["b"]["d"] is None
    spec = get_partition_spec(t)
    assert spec["a"] == P() and spec["b"]["c"] == P("data", None)
    paths = [p for p, _ in tree_leaves_with_path(t, is_leaf=lambda x: isinstance(x, Partitioned))]
    assert paths == ["a", "b/c"]


def test_rng_streams_distinct_and_dropout_mask_rate():
    k = rng.PRNGKey(69)
    a, b = rng.split(k)
    assert rng.materialize(a) != rng.materialize(b)
    assert rng.materialize(rng.fold_in(a, 0)) != rng.materialize(rng.fold_in(a, 1))
    assert rng.materialize(rng.rebase(a)) == rng.mix64(rng.materialize(a) ^ rng.mix64(0))
    m = rng.dropout_mask(rng.materialize(a), 256, 512, 0.1)
    assert abs(m.float().mean().item() - 0.9) < 0.01
    m2 = rng.dropout_mask(rng.materialize(b), 256, 512, 0.1)
    assert (m != m2).float().mean() > 0.1


def test_adamw_matches_optax_formula():
    torch.manual_seed(0)
    p = {"w": torch.randn(5, 3), "b": torch.randn(3)}
    tx = adamw(1e-3)
    st = tx.init(p)
    ref_p = {k: v.clone().double() for k, v in p.items()}
    mu = {k: torch.zeros_like(v).double() for k, v in p.items()}
    nu = {k: torch.zeros_like(v).double() for k, v in p.items()}
    for t in range(1, 5):
        g = {k: torch.randn_like(v) for k, v in p.items()}
        upd, st = tx.update(g, st, p)
        p = tx.apply_updates(p, upd)
        for k in ref_p:
            mu[k] = 0.9 * mu[k] + 0.1 * g[k].double()
            nu[k] = 0.999 * nu[k] + 0.001 * g[k].double() ** 2
            mh, vh = mu[k] / (1 - 0.9 ** t), nu[k] / (1 - 0.999 ** t)
            ref_p[k] = ref_p[k] - 1e-3 * (mh / (vh.sqrt() + 1e-8) + 1e-4 * ref_p[k])
    for k in p:
        assert torch.allclose(p[k].double(), ref_p[k], atol=1e-6)
    assert st[0].count == 4
    # optax state layout for checkpoints
    names = [path for path, _ in tree_leaves_with_path(st)]
    assert "0/count" in names and "0/mu/w" in names and "0/nu/b" in names


def _make_state(cfg, mesh):
    model = Classifier(cfg.model)
    key = rng.PRNGKey(cfg.seed)
    model_rng, data_rng, label_rng = rng.split(key, 3)
    g = rng.uniform_generator(data_rng)
    inputs = torch.randn(cfg.data.batch_size, cfg.data.input_size, generator=g)
    labels = torch.randint(0, cfg.data.num_classes, (cfg.data.batch_size,), generator=rng.uniform_generator(label_rng))
    with mesh:
        state = dp.init_dp(model_rng, inputs, model)
    return model, state, Batch(inputs, labels)


def test_classifier_param_tree_and_count():
    cfg = dp_config()
    mesh = Mesh((1,), ("data",))
    model, state, batch = _make_state(cfg, mesh)
    assert set(state.params) == {"input_dense", "output_dense"}
    assert tuple(state.params["input_dense"]["kernel"].shape) == (784, 512)
    assert tuple(state.params["output_dense"]["kernel"].shape) == (512, 10)
    assert state.params["input_dense"]["kernel"].dtype == torch.float32
    assert get_num_params(state) == 407050
    std = state.params["input_dense"]["kernel"].std().item()
    assert abs(std - math.sqrt(1 / 784)) < 0.003
    logits = model.apply({"params": state.params}, batch.inputs, train=False)
    assert logits.shape == (128, 10) and logits.dtype == torch.float32


def test_accum_loop_equals_scan_and_full_batch():
    cfg = dp_config()
    cfg.model.dropout_rate = 0.0
    mesh = Mesh((1,), ("data",))
    model, state, batch = _make_state(cfg, mesh)
    with mesh:
        key = rng.PRNGKey(1)
        g_loop, m_loop = accum_grads(state, batch, key, 4, dp.loss_fn, use_scan=False)
        g_scan, m_scan = accum_grads(state, batch, key, 4, dp.loss_fn, use_scan=True)
        g_full, m_full = accum_grads(state, batch, key, 1, dp.loss_fn)
    for a, b, c in zip(tree_leaves(g_loop), tree_leaves(g_scan), tree_leaves(g_full)):
        assert torch.allclose(a, b, atol=1e-6)
        assert torch.allclose(a, c, atol=2e-3)   # bf16 matmuls, different batch tiling
    assert m_loop["loss"][1] == 128 and m_loop["accuracy"][1] == 128      # metrics are summed (util.py:75)
    assert torch.allclose(m_loop["loss"][0], m_scan["loss"][0]) and torch.allclose(m_loop["loss"][0], m_full["loss"][0], rtol=1e-3)


def test_train_step_dp_overfits_single_batch(capsys):
    cfg = dp_config()
    mesh = Mesh((1,), ("data",))
    model, state, batch = _make_state(cfg, mesh)
    step = dp.make_train_step_dp(mesh, num_minibatches=cfg.optimizer.num_minibatches)
    metrics = None
    first = None
    for i in range(11):
        state, m = step(state, None, batch)
        if first is None:
            first = m["loss"][0].item() / m["loss"][1]
    last = m["loss"][0].item() / m["loss"][1]
    assert last < first * 0.7, (first, last)
    assert state.step == 11
    s = format_metrics(m, "dp")
    assert s.splitlines()[0].strip("=") == " dp " and "loss: " in s and "accuracy: " in s


def test_flat_layout_views_and_owner_slices_round_trip():
    """The engines' flat padded layout (host-side index math only, no GPU): leaf views address the
    right elements, pads stay out of them, and the owner-update slicing of data parallelism
    (each rank keeps 1/world of every bucket) reassembles to the flat buffer for any world size."""
    from jax_distributed_tuts_b200.engine.base import FlatLayout
    leaves = [("out/kernel", (512, 10), 0), ("in/kernel", (784, 512), 1), ("in/bias", (512,), 1), ("out/bias", (10,), 1)]
    for world in (1, 2, 4, 8):
        lay = FlatLayout(leaves, world)
        assert lay.total % (world * 64) == 0 and all((b1 - b0) % world == 0 for b0, b1 in lay.bucket_ranges)
        flat = torch.arange(lay.total, dtype=torch.float32)
        # leaf views: shapes, and the 10-wide leaf is stored with a pitch of 16 (TMA / vector alignment)
        assert lay.view(flat, "out/kernel").shape == (512, 10) and lay.leaves["out/kernel"].ld == 16
        assert lay.view(flat, "in/bias").shape == (512,)
        v = lay.view2d(flat, "out/kernel")
        assert v[1, 0] - v[0, 0] == 16 and v[0, 9] - v[0, 0] == 9
        # no two leaves overlap
        seen = torch.zeros(lay.total, dtype=torch.int32)
        for path, sp in lay.leaves.items():
            seen[sp.offset: sp.offset + sp.rows * sp.ld] += 1
        assert int(seen.max()) == 1
        # owner slices <-> flat
        slices = torch.stack([lay.flat_to_owner_slice(flat, r) for r in range(world)])
        assert slices.shape == (world, lay.total // world)
        assert torch.equal(lay.owner_slices_to_flat(slices), flat)
        # every flat element is owned by exactly one rank
        assert torch.equal(torch.sort(slices.reshape(-1)).values, flat)


def test_print_exception_names_the_class(capsys):
    from jax_distributed_tuts_b200.utils import print_exception
    try:
        raise ValueError("bad shape (3, 5)")
    except ValueError as e:
        print_exception(e)
    out = capsys.readouterr().out
    assert "ValueError" in out and "bad shape" in out


def test_sgd_matches_the_optax_formula_and_trains():
    """optax.sgd: trace_t = g + momentum * trace_{t-1}; update = -lr * trace_t (nesterov: -lr * (g + momentum * trace_t))."""
    import torch
    from jax_distributed_tuts_b200.optim import sgd
    p = {"w": torch.tensor([1.0, -2.0]), "b": torch.tensor([0.5])}
    g1 = {"w": torch.tensor([0.1, 0.2]), "b": torch.tensor([-1.0])}
    g2 = {"w": torch.tensor([-0.3, 0.0]), "b": torch.tensor([0.5])}
    for nesterov in (False, True):
        tx = sgd(0.1, momentum=0.9, nesterov=nesterov)
        st = tx.init(p)
        u1, st = tx.update(g1, st, p)
        p1 = tx.apply_updates(p, u1)
        u2, st = tx.update(g2, st, p1)
        t1 = g1["w"]
        t2 = g2["w"] + 0.9 * t1
        want2 = -0.1 * (g2["w"] + 0.9 * t2) if nesterov else -0.1 * t2
        assert torch.allclose(u2["w"], want2) and torch.allclose(st[0].trace["w"], t2)
    plain = sgd(0.5)
    u, st = plain.update(g1, plain.init(p), p)
    assert torch.allclose(u["b"], torch.tensor([0.5])) and len(st) == 2
    # plugs into the reference's train step
    from jax_distributed_tuts_b200 import rng
    from jax_distributed_tuts_b200.config import dp_config
    from jax_distributed_tuts_b200.mesh import Mesh
    from jax_distributed_tuts_b200.models.mlp import Classifier
    from jax_distributed_tuts_b200.parallel import dp
    from jax_distributed_tuts_b200.state import Batch
    cfg = dp_config()
    cfg.model.dropout_rate = 0.0
    mesh = Mesh((1,), ("data",), device=torch.device("cpu"))
    x = torch.randn(32, 784, generator=torch.Generator().manual_seed(0))
    y = torch.randint(0, 10, (32,), generator=torch.Generator().manual_seed(1))
    with mesh:
        state = dp.init_dp(rng.PRNGKey(0), x, Classifier(cfg.model), optimizer=sgd(0.05, momentum=0.9))
        step = dp.make_train_step_dp(mesh, num_minibatches=2)
        losses = []
        for _ in range(8):
            state, m = step(state, None, Batch(x, y))
            losses.append(float(m["loss"][0]) / m["loss"][1])
    assert losses[-1] < 0.7 * losses[0], losses


def test_learning_rate_schedules():
    import torch
    from jax_distributed_tuts_b200.optim import adamw, sgd, warmup_cosine_decay_schedule, linear_schedule
    s = warmup_cosine_decay_schedule(0.0, 1e-3, warmup_steps=10, decay_steps=100, end_value=1e-4)
    assert s(0) == 0.0 and abs(s(5) - 5e-4) < 1e-12 and abs(s(10) - 1e-3) < 1e-12
    assert abs(s(100) - 1e-4) < 1e-9 and abs(s(1000) - 1e-4) < 1e-9 and s(30) > s(60) > s(90)
    lin = linear_schedule(1.0, 0.0, 4, transition_begin=2)
    assert [lin(c) for c in (0, 2, 4, 6, 9)] == [1.0, 1.0, 0.5, 0.0, 0.0]
    # the optimizers evaluate the schedule at the number of updates done so far
    p = {"w": torch.ones(2)}
    g = {"w": torch.ones(2)}
    tx = sgd(lin)
    st = tx.init(p)
    seen = []
    for _ in range(5):
        u, st = tx.update(g, st, p)
        seen.append(-u["w"][0].item())
    assert seen == [1.0, 1.0, 1.0, 0.75, 0.5] and st[1].count == 5
    ta = adamw(lambda count: 0.1 if count == 0 else 0.0, weight_decay=0.0)
    sa = ta.init(p)
    u0, sa = ta.update(g, sa, p)
    u1, sa = ta.update(g, sa, p)
    assert abs(u0["w"][0].item() + 0.1) < 1e-6 and u1["w"][0].item() == 0.0

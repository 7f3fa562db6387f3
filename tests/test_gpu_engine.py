"""GPU tier: generic autograd path (our kernels behind torch.autograd) vs the CPU oracle, and
the fused CUDA-graph engine vs the generic path, on the reference's MLP."""
import pytest
import torch

from jax_distributed_tuts_b200 import rng
from jax_distributed_tuts_b200.config import dp_config
from jax_distributed_tuts_b200.mesh import Mesh
from jax_distributed_tuts_b200.models.mlp import Classifier
from jax_distributed_tuts_b200.parallel import dp
from jax_distributed_tuts_b200.state import Batch
from jax_distributed_tuts_b200.tree import tree_leaves, tree_map

pytestmark = pytest.mark.gpu


def _setup(device, batch=64, dropout=0.1):
    cfg = dp_config()
    cfg.model.dropout_rate = dropout
    mesh = Mesh((1,), ("data",), device=torch.device(device))
    model = Classifier(cfg.model)
    key = rng.PRNGKey(cfg.seed)
    model_rng, data_rng, label_rng = rng.split(key, 3)
    x = torch.randn(batch, 784, generator=rng.uniform_generator(data_rng))
    y = torch.randint(0, 10, (batch,), generator=rng.uniform_generator(label_rng))
    with mesh:
        state = dp.init_dp(model_rng, x, model)      # CPU init: identical values for every path
    return cfg, mesh, model, state, Batch(x, y)


def _to(state, device):
    return state.replace(params=tree_map(lambda t: t.to(device), state.params),
                         opt_state=state.tx.init(tree_map(lambda t: t.to(device), state.params)),
                         rng=state.rng.to(device))


def _mean_close(a, b, lr, steps):
    d = (a.float().cpu() - b.float().cpu()).abs()
    return d.max() <= 2.2 * lr * steps and d.mean() <= 0.06 * lr * steps


def test_generic_gpu_path_matches_cpu_oracle_with_dropout():
    cfg, mesh, model, state, batch = _setup("cpu")
    steps, lr = 3, cfg.optimizer.learning_rate
    step = dp.make_train_step_dp(mesh, num_minibatches=4)
    s_cpu, m_cpu = state, None
    for _ in range(steps):
        s_cpu, m_cpu = step(s_cpu, m_cpu, batch)
    mesh_g = Mesh((1,), ("data",), device=torch.device("cuda"))
    step_g = dp.make_train_step_dp(mesh_g, num_minibatches=4)
    s_gpu, m_gpu = _to(state, "cuda"), None
    bg = batch.to("cuda")
    for _ in range(steps):
        s_gpu, m_gpu = step_g(s_gpu, m_gpu, bg)
    for a, b in zip(tree_leaves(s_cpu.params), tree_leaves(s_gpu.params)):
        assert _mean_close(a, b, lr, steps), (a.cpu() - b.cpu()).abs().mean()
    l_cpu = float(m_cpu["loss"][0]) / m_cpu["loss"][1]
    l_gpu = float(m_gpu["loss"][0]) / m_gpu["loss"][1]
    assert abs(l_cpu - l_gpu) < 2e-2, (l_cpu, l_gpu)
    assert abs(float(m_cpu["accuracy"][0]) - float(m_gpu["accuracy"][0])) <= 3


@pytest.mark.parametrize("use_graph,fuse", [(False, False), (False, True), (True, True)])
def test_fused_engine_matches_generic_path(use_graph, fuse):
    from jax_distributed_tuts_b200.engine import MLPEngine
    cfg, _, model, state, batch = _setup("cpu")
    steps, lr = 4, cfg.optimizer.learning_rate
    mesh_g = Mesh((1,), ("data",), device=torch.device("cuda"))
    step = dp.make_train_step_dp(mesh_g, num_minibatches=4)
    s_gen, m_gen = _to(state, "cuda"), None
    bg = batch.to("cuda")
    for _ in range(steps):
        s_gen, m_gen = step(s_gen, m_gen, bg)
    with mesh_g:
        eng = MLPEngine(cfg.model, mesh_g, mode="dp", batch_per_rank=batch.inputs.shape[0], lr=lr,
                        num_minibatches=4, init_params=state.params, use_graph=use_graph,
                        fuse_minibatches=fuse)
        s_eng = eng.make_state(model, state.rng)
        m_eng = None
        pinned = Batch(batch.inputs.pin_memory(), batch.labels.pin_memory())
        for _ in range(steps):
            s_eng, m_eng = step(s_eng, m_eng, pinned)
    torch.cuda.synchronize()
    p_eng = eng.params_tree()
    for k in ("input_dense", "output_dense"):
        for kk in ("kernel", "bias"):
            assert _mean_close(p_eng[k][kk], s_gen.params[k][kk], lr, steps), (k, kk)
    l_gen = float(m_gen["loss"][0]) / m_gen["loss"][1]
    l_eng = float(m_eng["loss"][0]) / m_eng["loss"][1]
    assert m_eng["loss"][1] == m_gen["loss"][1] == steps * batch.inputs.shape[0]
    assert abs(l_gen - l_eng) < 2e-2, (l_gen, l_eng)
    assert s_eng.step == steps and int(eng.step_count.item()) == steps
    assert eng.launches_per_step > 0


def test_engine_fsdp_single_rank_matches_dp_engine():
    from jax_distributed_tuts_b200.engine import MLPEngine
    cfg, _, model, state, batch = _setup("cpu", dropout=0.1)
    mesh_g = Mesh((1,), ("data",), device=torch.device("cuda"))
    res = {}
    with mesh_g:
        for mode in ("dp", "fsdp"):
            eng = MLPEngine(cfg.model, mesh_g, mode=mode, batch_per_rank=64, lr=1e-3, num_minibatches=4,
                            init_params=state.params, use_graph=False, min_weight_size=16)
            st = eng.make_state(model, state.rng)
            m = None
            for _ in range(3):
                st, m = eng.train_step(st, m, batch)
            torch.cuda.synchronize()
            res[mode] = (eng.params_tree(), float(m["loss"][0]))
    for k in ("input_dense", "output_dense"):
        for kk in ("kernel", "bias"):
            a, b = res["dp"][0][k][kk], res["fsdp"][0][k][kk]
            assert torch.allclose(a.float(), b.float().reshape(a.shape), atol=1e-6), (k, kk)
    assert abs(res["dp"][1] - res["fsdp"][1]) < 1e-4


def test_smoke_entry_point():
    import __graft_entry__ as g
    g.smoke()


def _gpt_setup(name="tiny", batch=4, seq=128):
    from jax_distributed_tuts_b200.models.gpt2 import GPT2, gpt2_config
    cfg = gpt2_config(name)
    model = GPT2(cfg)
    g = torch.Generator().manual_seed(0)
    tok = torch.randint(0, cfg.vocab_size, (batch, seq + 1), generator=g)
    params = model.init({"params": rng.PRNGKey(3)}, tok[:, :seq])["params"]
    return cfg, model, params, Batch(tok[:, :seq].contiguous(), tok[:, 1:].contiguous())


def _leaves_with_path(tree):
    from jax_distributed_tuts_b200.tree import tree_leaves_with_path
    return tree_leaves_with_path(tree)


def test_gpt2_generic_gpu_matches_cpu_oracle():
    from jax_distributed_tuts_b200.autodiff import value_and_grad
    from jax_distributed_tuts_b200.models.gpt2 import lm_loss_fn
    cfg, model, params, batch = _gpt_setup()
    vg = value_and_grad(lm_loss_fn, has_aux=True)
    (l_cpu, _), g_cpu = vg(params, model.apply, batch, rng.PRNGKey(1))
    pg = tree_map(lambda t: t.cuda(), params)
    (l_gpu, m_gpu), g_gpu = vg(pg, model.apply, batch.to("cuda"), rng.PRNGKey(1))
    assert abs(l_cpu.item() - l_gpu.item()) < 2e-2, (l_cpu.item(), l_gpu.item())
    for (pa, a), (pb, b) in zip(_leaves_with_path(g_cpu), _leaves_with_path(g_gpu)):
        denom = a.abs().max().clamp_min(1e-6)
        rel = ((a - b.cpu().float()).abs().max() / denom).item()
        assert rel < 6e-2, (pa, rel)


@pytest.mark.parametrize("nmb", [1, 2])
def test_gpt2_engine_matches_generic_path(nmb):
    from jax_distributed_tuts_b200.engine import GPT2Engine
    from jax_distributed_tuts_b200.models.gpt2 import lm_loss_fn
    from jax_distributed_tuts_b200.optim import adamw
    from jax_distributed_tuts_b200.state import TrainState
    cfg, model, params, batch = _gpt_setup()
    lr, steps = 1e-3, 3
    mesh_g = Mesh((1,), ("data",), device=torch.device("cuda"))
    pg = tree_map(lambda t: t.cuda(), params)
    state = TrainState.create(apply_fn=model.apply, params=pg, tx=adamw(lr), rng=rng.PRNGKey(5).to("cuda"))
    step = dp.make_train_step_dp(mesh_g, num_minibatches=nmb, loss_fn=lm_loss_fn)
    m_gen, bg = None, batch.to("cuda")
    losses_gen = []
    for _ in range(steps):
        state, m_gen = step(state, None, bg)
        losses_gen.append(float(m_gen["loss"][0]) / m_gen["loss"][1])
    with mesh_g:
        eng = GPT2Engine(cfg, mesh_g, mode="dp", batch_per_rank=batch.inputs.shape[0], seq_len=128, lr=lr,
                         num_minibatches=nmb, init_params=params, use_graph=False)
        st = eng.make_state(model, rng.PRNGKey(5))
        losses_eng = []
        for _ in range(steps):
            st, m = eng.train_step(st, None, batch)
            losses_eng.append(float(m["loss"][0]) / m["loss"][1])
    torch.cuda.synchronize()
    assert m["loss"][1] == batch.labels.numel()
    for a, b in zip(losses_gen, losses_eng):
        assert abs(a - b) < 3e-2, (losses_gen, losses_eng)
    assert losses_eng[-1] < losses_eng[0]
    p_eng = eng.params_tree()
    for (pa, a), (pb, b) in zip(_leaves_with_path(state.params), _leaves_with_path(p_eng)):
        assert pa == pb
        bb = b[: a.shape[0]] if b.shape != a.shape else b
        d = (a.float().cpu() - bb.float().cpu()).abs()
        # Adam turns rounding noise on (near-)zero gradients into +-lr steps; the key bias of
        # attention has an exactly zero gradient (softmax is shift invariant), so bound biases loosely
        mean_tol = (0.6 if pa.endswith("bias") else 0.06) * lr * steps
        assert d.max() <= 2.2 * lr * steps and d.mean() <= mean_tol, (pa, d.max(), d.mean())


@pytest.mark.parametrize("ext", ["msgpack", "safetensors"])
def test_engine_checkpoint_has_the_generic_layout_and_resumes(tmp_path, ext):
    """An engine checkpoint uses the same container as a generic-tier TrainState (mu / nu shaped like
    the parameters, global layout), and a fresh engine restored from it continues like the original."""
    from jax_distributed_tuts_b200 import checkpoint as ckpt
    from jax_distributed_tuts_b200.engine import MLPEngine
    cfg, _, model, state, batch = _setup("cpu")
    mesh = Mesh((1,), ("data",), device=torch.device("cuda"))
    lr = cfg.optimizer.learning_rate
    pinned = Batch(batch.inputs.pin_memory(), batch.labels.pin_memory())
    with mesh:
        eng = MLPEngine(cfg.model, mesh, mode="dp", batch_per_rank=batch.inputs.shape[0], lr=lr, num_minibatches=4,
                        init_params=state.params, use_graph=True)
        st = eng.make_state(model, state.rng)
        for _ in range(3):
            st, _m = eng.train_step(st, None, pinned)
        path = str(tmp_path / f"ckpt.{ext}")
        ckpt.save_checkpoint(path, st, mesh=mesh)
        loaded = ckpt.load_checkpoint(path)
        # container: identical key structure to the generic tier's
        generic = ckpt.state_dict(_to(state, "cuda"), mesh=mesh)
        def keys(t, p=""):
            return sorted(k2 for k, v in t.items() for k2 in (keys(v, p + k + "/") if isinstance(v, dict) else [p + k]))
        assert keys(loaded["params"]) == keys(generic["params"])
        assert keys(loaded["opt_state"]["0"]) == keys(generic["opt_state"]["0"])
        assert int(loaded["opt_state"]["0"]["count"]) == 3
        eng2 = MLPEngine(cfg.model, mesh, mode="dp", batch_per_rank=batch.inputs.shape[0], lr=lr, num_minibatches=4,
                         init_params=tree_map(torch.zeros_like, state.params), use_graph=False)
        st2 = eng2.make_state(model, state.rng)
        ckpt.restore_engine(eng2, loaded)
        st, m1 = eng.train_step(st, None, pinned)
        st2, m2 = eng2.train_step(st2, None, pinned)
        torch.cuda.synchronize()
        for a, b in zip(tree_leaves(eng.params_tree()), tree_leaves(eng2.params_tree())):
            assert (a.float() - b.float()).abs().max().item() <= 2e-5
        assert abs(float(m1["loss"][0]) - float(m2["loss"][0])) <= 1e-2 * max(1.0, abs(float(m1["loss"][0])))


@pytest.mark.parametrize("num_classes,fuse_xent", [(10, False), (40, True)])
def test_engine_head_variants_match_each_other(num_classes, fuse_xent):
    """The classifier head runs either with the cross-entropy fused into the GEMM epilogue (<= 32
    classes) or as GEMM + cross-entropy kernel (more classes, or fuse_xent=False): both must follow
    the generic path."""
    from jax_distributed_tuts_b200.engine import MLPEngine
    cfg = dp_config()
    cfg.model.num_classes = num_classes
    mesh_g = Mesh((1,), ("data",), device=torch.device("cuda"))
    model = Classifier(cfg.model)
    key = rng.PRNGKey(cfg.seed)
    model_rng, data_rng, label_rng = rng.split(key, 3)
    x = torch.randn(64, 784, generator=rng.uniform_generator(data_rng))
    y = torch.randint(0, num_classes, (64,), generator=rng.uniform_generator(label_rng))
    batch = Batch(x, y)
    with Mesh((1,), ("data",), device=torch.device("cpu")):
        state = dp.init_dp(model_rng, x, model)
    steps, lr = 3, cfg.optimizer.learning_rate
    step = dp.make_train_step_dp(mesh_g, num_minibatches=4)
    s_gen, m_gen = _to(state, "cuda"), None
    bg = batch.to("cuda")
    for _ in range(steps):
        s_gen, m_gen = step(s_gen, m_gen, bg)
    with mesh_g:
        eng = MLPEngine(cfg.model, mesh_g, mode="dp", batch_per_rank=64, lr=lr, num_minibatches=4,
                        init_params=state.params, use_graph=False, fuse_xent=fuse_xent)
        s_eng, m_eng = eng.make_state(model, state.rng), None
        for _ in range(steps):
            s_eng, m_eng = eng.train_step(s_eng, m_eng, batch)
    torch.cuda.synchronize()
    assert eng.launches_per_step == 8        # neither variant takes the 7-launch fused-head path
    p_eng = eng.params_tree()
    for k in ("input_dense", "output_dense"):
        for kk in ("kernel", "bias"):
            assert _mean_close(p_eng[k][kk], s_gen.params[k][kk], lr, steps), (k, kk)
    assert abs(float(m_gen["loss"][0]) / m_gen["loss"][1] - float(m_eng["loss"][0]) / m_eng["loss"][1]) < 2e-2
    assert float(m_gen["accuracy"][0]) == pytest.approx(float(m_eng["accuracy"][0]), abs=3)


def test_accum_scan_is_a_captured_graph_and_equals_the_loop():
    """accum_grads(use_scan=True) on a GPU replays ONE captured minibatch body (util.py:81-137 compiles one
    body instead of n): same gradients as the loop with dropout off, and no per-minibatch kernel launches
    from the host besides the key write."""
    from jax_distributed_tuts_b200.accum import accum_grads, _SCAN_CACHE
    from jax_distributed_tuts_b200.ops import _lib
    cfg, _, model, state, batch = _setup("cpu", batch=64, dropout=0.0)
    mesh_g = Mesh((1,), ("data",), device=torch.device("cuda"))
    st, bg = _to(state, "cuda"), batch.to("cuda")
    key = rng.PRNGKey(3, device="cuda")
    with mesh_g:
        g_loop, m_loop = accum_grads(st, bg, key, 4, dp.loss_fn, use_scan=False)
        _SCAN_CACHE.clear()
        g_scan, m_scan = accum_grads(st, bg, key, 4, dp.loss_fn, use_scan=True)      # captures
        torch.cuda.synchronize()
        before = _lib.launch_count()
        g_scan2, m_scan2 = accum_grads(st, bg, key, 4, dp.loss_fn, use_scan=True)    # replays only
        torch.cuda.synchronize()
        launches = _lib.launch_count() - before
    assert len(_SCAN_CACHE) == 1 and next(iter(_SCAN_CACHE.values())).replays == 8
    assert launches <= 4, f"{launches} library launches for 4 replayed minibatches (only the key writes are expected)"
    for a, b, c in zip(tree_leaves(g_loop), tree_leaves(g_scan), tree_leaves(g_scan2)):
        assert torch.equal(a, b) and torch.equal(b, c)
    assert torch.equal(m_loop["loss"][0], m_scan["loss"][0]) and m_loop["loss"][1] == m_scan["loss"][1] == 64
    assert torch.equal(m_loop["accuracy"][0], m_scan2["accuracy"][0])


def test_engine_gradients_match_fp32_pytorch():
    """One fused MLP step with dropout off: the flat fp32 gradient buffer against autograd of the same model in
    plain fp32 PyTorch on the bf16-rounded operands, element-wise (atol + rtol * |ref|), before any optimizer
    arithmetic can blur a wrong sign or a missing term."""
    from jax_distributed_tuts_b200.engine import MLPEngine
    cfg, _, model, state, batch = _setup("cpu", batch=64, dropout=0.0)
    mesh_g = Mesh((1,), ("data",), device=torch.device("cuda"))
    with mesh_g:
        eng = MLPEngine(cfg.model, mesh_g, mode="dp", batch_per_rank=64, lr=1e-3, num_minibatches=4,
                        init_params=state.params, use_graph=False)
        s_eng = eng.make_state(model, state.rng)
        eng.train_step(s_eng, None, Batch(batch.inputs.pin_memory(), batch.labels.pin_memory()))
        torch.cuda.synchronize()
    bf = lambda t: t.to(torch.bfloat16).float()
    P = {k: {kk: bf(v).clone().requires_grad_(True) for kk, v in sub.items()} for k, sub in state.params.items()}
    x = bf(batch.inputs)
    total = 0.0
    for i in range(4):      # mean over minibatches of the minibatch-mean loss (util.py:58-77)
        xs, ys = x[i * 16:(i + 1) * 16], batch.labels[i * 16:(i + 1) * 16]
        pre = bf(xs @ P["input_dense"]["kernel"] + P["input_dense"]["bias"])
        h = bf(torch.nn.functional.silu(pre))
        logits = h @ P["output_dense"]["kernel"] + P["output_dense"]["bias"]
        total = total + torch.nn.functional.cross_entropy(logits, ys) / 4
    total.backward()
    for k in ("input_dense", "output_dense"):
        for kk in ("kernel", "bias"):
            got = eng.layout.view(eng.g32, f"{k}/{kk}").float().cpu() / 4      # the buffer holds the sum over minibatches
            ref = P[k][kk].grad
            err = (got - ref).abs()
            tol = 2e-4 + 3e-2 * ref.abs()
            assert bool((err <= tol).all()), f"{k}/{kk}: {int((err > tol).sum())} of {err.numel()} elements off, max {err.max():.3e}"


def test_engine_is_bitwise_reproducible():
    """Two runs of the same 3 steps with deterministic=True (split-K off, ordered bias-gradient sums) end in
    bit-identical parameters and Adam moments; the default schedule (arrival-order sums) agrees to rounding."""
    from jax_distributed_tuts_b200.engine import MLPEngine
    cfg, _, model, state, batch = _setup("cpu", batch=64, dropout=0.1)
    mesh_g = Mesh((1,), ("data",), device=torch.device("cuda"))
    pinned = Batch(batch.inputs.pin_memory(), batch.labels.pin_memory())

    def run(det):
        with mesh_g:
            eng = MLPEngine(cfg.model, mesh_g, mode="dp", batch_per_rank=64, lr=1e-3, num_minibatches=4,
                            init_params=state.params, deterministic=det)
            st = eng.make_state(model, state.rng)
            m = None
            for _ in range(3):
                st, m = eng.train_step(st, m, pinned)
            torch.cuda.synchronize()
            return [t.clone() for t in tree_leaves(eng.params_tree())] + [eng.m.clone(), eng.v.clone()], float(m["loss"][0])

    a, la = run(True)
    b, lb = run(True)
    c, lc = run(False)
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    assert abs(la - lb) < 1e-3 and abs(la - lc) < 1e-2
    for x, z in zip(a[:4], c[:4]):     # (Adam's first steps move by lr * sign(g): a rounding-level gradient can flip an element)
        assert (x - z).abs().mean().item() < 1e-5

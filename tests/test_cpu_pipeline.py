"""Pipeline parallelism on the CPU plumbing tier: GPipe over 2 and 4 stages == the unsplit model."""
import torch

from jax_distributed_tuts_b200.launch import sim_multiCPU_dev
from jax_distributed_tuts_b200.parallel.pp import split_layers

import _workers as W


def test_split_layers():
    assert [list(r) for r in split_layers(12, 4)] == [[0, 1, 2], [3, 4, 5], [6, 7, 8], [9, 10, 11]]
    assert [len(r) for r in split_layers(10, 4)] == [3, 3, 2, 2]
    assert [list(r) for r in split_layers(2, 1)] == [[0, 1]]


def _merge(results):
    """Union of the stages' parameter dicts (data replica 0)."""
    out = {}
    for r in results:
        if r["data"] == 0:
            for k, v in r["params"].items():
                out.setdefault(k, v)
    return out


def _close(a, b, lr=1e-3, steps=2, bias=False):
    d = (a.float() - b.float()).abs()
    return d.max() <= 2.2 * lr * steps and d.mean() <= (0.6 if bias else 0.06) * lr * steps


def _flat(tree, prefix=""):
    for k, v in tree.items():
        if isinstance(v, dict):
            yield from _flat(v, f"{prefix}{k}/")
        else:
            yield f"{prefix}{k}", v


def test_gpipe_matches_unsplit_model():
    steps, mb = 2, 2
    single = sim_multiCPU_dev(W.pp_worker, 1, args=(1, steps, mb))
    ref = dict(_flat(_merge(single)))
    for n_stages in (2, 4):
        res = sim_multiCPU_dev(W.pp_worker, n_stages, args=(n_stages, steps, mb))
        # every stage reports the same global metrics
        l0 = res[0]["metrics"]["loss"]
        assert l0[1] == single[0]["metrics"]["loss"][1] == steps * 4 * 128
        assert abs(float(l0[0]) - float(single[0]["metrics"]["loss"][0])) < 0.02 * l0[1]
        for r in res[1:]:
            assert float(r["metrics"]["loss"][0]) == float(l0[0])
        # first and last stage hold identical copies of the tied embedding
        assert torch.equal(res[0]["params"]["wte"], res[-1]["params"]["wte"])
        got = dict(_flat(_merge(res)))
        assert set(got) == set(ref)
        for k in ref:
            assert _close(got[k], ref[k], bias=k.endswith("bias")), (n_stages, k, (got[k] - ref[k]).abs().mean())


def test_gpipe_with_data_axis():
    res = sim_multiCPU_dev(W.pp_worker, 4, args=(2, 1, 2))      # mesh (data=2, pipe=2)
    by = {(r["data"], r["stage"]): r for r in res}
    for stage in (0, 1):   # data replicas of a stage stay identical
        a, b = dict(_flat(by[(0, stage)]["params"])), dict(_flat(by[(1, stage)]["params"]))
        for k in a:
            assert torch.equal(a[k], b[k]), (stage, k)
    assert by[(0, 0)]["metrics"]["loss"][1] == 2 * 4 * 128


def test_pipeline_checkpoint_holds_the_whole_model(tmp_path):
    """A pipeline run saves ONE global checkpoint (stages merged over the pipe axis, tied wte once) and each
    stage restores its own blocks from it; saving a single stage's state without naming the axis is refused."""
    path = str(tmp_path / "pp.msgpack")
    res = sim_multiCPU_dev(W.pp_ckpt_worker, 2, args=(path,))
    want = ["h_0", "h_1", "h_2", "h_3", "ln_f", "wpe", "wte"]
    for r in res:
        assert r["refused_without_axis"] and r["reload_equal"]
        assert r["top_keys"] == want and r["mu_keys"] == want


def test_balanced_split_puts_fewer_blocks_next_to_the_head():
    from jax_distributed_tuts_b200.engine.pp import balanced_split
    assert [len(r) for r in balanced_split(12, 4, 5.46)] == [4, 4, 4, 0]
    assert [len(r) for r in balanced_split(24, 4, 4.1)] == [7, 7, 7, 3]
    assert [len(r) for r in balanced_split(12, 1, 5.46)] == [12]
    s = balanced_split(4, 4, 0.3)
    assert [list(r) for r in s] == [[0], [1], [2], [3]]


def _simulate(schedule, S, M):
    """Runs the stages' sequences against each other the way the engine does: a stage's next operation
    needs its input message (forward: activations from below, backward: gradients from above), messages
    between a pair and direction are FIFO (one communicator each, receives posted in the receiver's own
    order), sends never block. Returns True when every stage finishes, i.e. no deadlock and every
    receive meets the send its peer issues at that position."""
    from jax_distributed_tuts_b200.engine.pp import stage_sequence
    seqs = [stage_sequence(schedule, S, s, M) for s in range(S)]
    pos = [0] * S
    up = [[] for _ in range(S)]      # up[s]: activations sent by stage s-1 to s, in send order
    down = [[] for _ in range(S)]    # down[s]: gradients sent by stage s+1 to s
    took_up, took_down = [0] * S, [0] * S
    progress = True
    while progress:
        progress = False
        for s in range(S):
            if pos[s] == len(seqs[s]):
                continue
            kind, k = seqs[s][pos[s]]
            if kind == "F":
                if s > 0:
                    if took_up[s] >= len(up[s]):
                        continue
                    assert up[s][took_up[s]] == k, (schedule, S, M, s, "forward receive order")
                    took_up[s] += 1
                if s + 1 < S:
                    up[s + 1].append(k)
            else:
                if s + 1 < S:
                    if took_down[s] >= len(down[s]):
                        continue
                    assert down[s][took_down[s]] == k, (schedule, S, M, s, "backward receive order")
                    took_down[s] += 1
                if s > 0:
                    down[s - 1].append(k)
            pos[s] += 1
            progress = True
    return all(p == len(q) for p, q in zip(pos, seqs))


def test_pipeline_schedules_are_deadlock_free_and_cover_every_microbatch():
    from jax_distributed_tuts_b200.engine.pp import stage_sequence
    for schedule in ("gpipe", "1f1b"):
        for S in range(1, 9):
            for M in (1, 2, 3, 4, 7, 16):
                for s in range(S):
                    seq = stage_sequence(schedule, S, s, M)
                    assert sorted(k for kind, k in seq if kind == "F") == list(range(M))
                    assert sorted(k for kind, k in seq if kind == "B") == list(range(M))
                    for k in range(M):      # a microbatch's backward never precedes its forward
                        assert seq.index(("F", k)) < seq.index(("B", k))
                assert _simulate(schedule, S, M), (schedule, S, M)

"""The block-to-segment plan of the fused update kernel (csrc/comm.cu: `segment_blocks`, the pass loop in
`fused_reduce_adamw_body`, the grid-stride loop of `update_segment`), restated in Python and checked by
property: for any segment list and grid, every 16-byte vector of every segment is handled by exactly one
(block, thread, unroll slot). The device code is exercised on GPUs by tools/gpu_check_comm.py; this mirror
guards the integer logic (runs of small segments side by side, large ones on the whole grid)."""
from hypothesis import given, settings, strategies as st

THREADS, U = 256, 4


def segment_blocks(elems: int) -> int:
    per_block = THREADS * 4 * U
    return max(1, min(1 << 20, (elems + per_block - 1) // per_block))


def plan(seg_elems, nblocks):
    """-> list of passes; a pass is a list of (segment, first_block, n_blocks)."""
    small_limit = max(1, nblocks // 4)
    passes, s = [], 0
    n = len(seg_elems)
    while s < n:
        need = segment_blocks(seg_elems[s])
        e = s + 1
        if need <= small_limit and s + 1 < n:
            run, first = [(s, 0, need)], need
            while e < n:
                n2 = segment_blocks(seg_elems[e])
                if n2 > small_limit or first + n2 > nblocks:
                    break
                run.append((e, first, n2))
                first += n2
                e += 1
            passes.append(run if e > s + 1 else [(s, 0, nblocks)])
        else:
            passes.append([(s, 0, nblocks)])
        s = e
    return passes


def covered(total4, vnblocks):
    """Vector indices touched by a (sub-)grid of vnblocks blocks running update_segment's loop."""
    nthreads = vnblocks * THREADS
    seen = []
    for tid in range(nthreads):
        i0 = tid
        while i0 < total4:
            for u in range(U):
                i = i0 + u * nthreads
                if i < total4:
                    seen.append(i)
            i0 += nthreads * U
    return seen


@settings(max_examples=60, deadline=None)
@given(st.lists(st.integers(min_value=1, max_value=3000), min_size=1, max_size=12), st.sampled_from([1, 2, 4, 8, 16]))
def test_every_vector_is_updated_exactly_once(vecs_per_seg, nblocks):
    seg_elems = [4 * v for v in vecs_per_seg]
    passes = plan(seg_elems, nblocks)
    assert sorted(s for run in passes for (s, _, _) in run) == list(range(len(seg_elems)))     # each segment in one pass
    for run in passes:
        used = 0
        for (s, first, nb) in run:
            assert first == used and nb >= 1            # disjoint, contiguous block ranges
            used += nb
            seen = covered(vecs_per_seg[s], nb)
            assert sorted(seen) == list(range(vecs_per_seg[s]))
        assert used <= nblocks


def test_plan_examples():
    # an MLP's four small leaves on 128 blocks: one pass, side by side
    assert plan([50176, 64, 1024, 16], 128) == [[(0, 0, 13), (1, 13, 1), (2, 14, 1), (3, 15, 1)]]
    # a large matrix takes the whole grid, the small vectors behind it share a pass
    p = plan([4_800_000, 768, 768, 3072], 128)
    assert p[0] == [(0, 0, 128)] and p[1] == [(1, 0, 1), (2, 1, 1), (3, 2, 1)]
    # a single small segment gets the whole grid (nothing to share it with)
    assert plan([4096], 128) == [[(0, 0, 128)]]

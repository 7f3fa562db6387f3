"""Allocation-free attention backward (``b200_attention_bwd_noalloc``), called through the C ABI.

Some users of the kernel library cannot give it scratch memory: a C host or plugin with no
device allocator, or a runtime that fixes its device-memory plan at start-up and bans
cudaMalloc afterwards. For them the library keeps the scratch in a built-in device array. This
test drives that entry point with ctypes exactly as such a host would (raw pointers, no torch
ops inside the call) and checks it against the regular entry point, which takes caller-provided
scratch, on two consecutive calls with different data - the second call meets an array that is
still full of the first call's numbers.

The file name sorts last on purpose: if it fails under ``-x`` every other test has already run.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ROWS = 16384  # B200_ATTN_BWD_NOALLOC_MAX_ROWS in include/b200dist.h


def _lib():
    from jax_distributed_tuts_b200.ops import _lib
    lib = _lib.lib()
    fn = lib.b200_attention_bwd_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p] * 5 + [C.c_int] * 3 + [C.c_void_p]
    return _lib, lib, fn


def _case(seed, b, t, h):
    from jax_distributed_tuts_b200.ops import kernels as K
    g = torch.Generator(device="cuda").manual_seed(seed)
    qkv = torch.randn(b, t, 3, h, 64, device="cuda", generator=g).bfloat16()
    dout = torch.randn(b, t, h, 64, device="cuda", generator=g).bfloat16()
    out = torch.empty(b, t, h, 64, device="cuda", dtype=torch.bfloat16)
    lse = torch.empty(b, h, t, device="cuda")
    K.attention_fwd(qkv, out, lse, b, t, h)
    return qkv, out, dout, lse


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-6)).item()


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    from jax_distributed_tuts_b200.ops import kernels as K
    _l, lib, noalloc = _lib()
    stream = torch.cuda.current_stream().cuda_stream
    for seed, (b, t, h) in ((11, (1, 256, 4)), (12, (2, 512, 3))):   # 1024 and 3072 rows, both within the limit
        qkv, out, dout, lse = _case(seed, b, t, h)
        ref = torch.empty_like(qkv)
        scratch = torch.empty(K.attention_bwd_scratch_numel(b, t, h), device="cuda")
        K.attention_bwd(qkv, out, dout, lse, ref, scratch, b, t, h)
        got = torch.full_like(qkv, float("nan"))
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = noalloc(qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), got.data_ptr(), b, t, h, stream)
        assert rc == 0, lib.b200_last_error().decode()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before       # nothing was taken from torch's allocator either
        assert not torch.isnan(got.float()).any()
        # same kernels, same inputs: only the order of the fp32 dQ additions may differ
        for i, name in enumerate(("dq", "dk", "dv")):
            assert _rel(got[:, :, i], ref[:, :, i]) <= 2e-2, name


def test_noalloc_rejects_problems_beyond_the_builtin_scratch():
    _l, lib, noalloc = _lib()
    b, t, h = 2, 1024, 12            # 24576 rows > MAX_ROWS
    assert b * t * h > MAX_ROWS
    x = torch.zeros(8, device="cuda")
    rc = noalloc(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), b, t, h,
                 torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.b200_last_error().decode()
    assert "exceeds the built-in scratch" in msg and str(MAX_ROWS) in msg

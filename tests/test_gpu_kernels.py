"""Single-GPU kernel tier: every sm_100a kernel against a plain PyTorch fp32 reference.
The case lists live in tools/gpu_check_kernels.py (also runnable stand-alone on a GPU box)."""
import pytest

from tools import gpu_check_kernels as G

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("group", ["gemm_kmajor", "gemm_mnmajor", "gemm_epilogue", "gemm_xent", "gemm_sharded", "elementwise", "transformer", "attention"])
def test_kernel_group(group):
    import torch
    assert getattr(G, "group_" + group)()
    torch.cuda.synchronize()


def test_cross_rank_protocol_emulated_on_one_gpu():
    """Fused reduce+AdamW+publish and the optimizer-free collectives with 2/4/8 virtual ranks in one
    cooperative launch: epoch flags, split-phase publish counters, metric mailboxes, dead-peer abort."""
    from tools import gpu_check_comm
    assert gpu_check_comm.group_comm_emulated()


def test_library_is_the_native_one():
    from jax_distributed_tuts_b200.ops import _lib
    before = _lib.launch_count()
    import torch
    from jax_distributed_tuts_b200.ops import kernels as K
    x = torch.ones(1024, device="cuda")
    K.scale_(x, 2.0)
    torch.cuda.synchronize()
    assert _lib.launch_count() == before + 1 and x[0].item() == 2.0


def test_no_kernel_writes_outside_its_outputs():
    """Every kernel family on odd shapes with 4 KiB guard bands around every tensor (the stand-in for a memory
    checker on this pool, see tools/smoke_all_kernels.py)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "smoke_all_kernels.py"), "--guard"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "0 overrun(s)" in r.stdout, (r.stdout + r.stderr)[-2000:]

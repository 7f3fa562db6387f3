"""Multi-GPU tier (needs >= 2 GPUs on the box; skipped otherwise): the fused collective kernels
against NCCL results. Each check is a torchrun job (one rank per GPU) that exits non-zero on any
mismatch; see tools/gpu_check_multi.py and tools/gpu_check_multi_gpt.py."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _torchrun(script, n, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tools", script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    tail = "\n".join((r.stdout + r.stderr).splitlines()[-30:])
    assert r.returncode == 0, tail
    return r.stdout


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
def test_fused_mlp_engines_match_nccl():
    out = _torchrun("gpu_check_multi.py", min(_ngpu(), 8), 29531)
    assert "MULTI-GPU CHECK PASS" in out


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
def test_fused_gpt2_engines_match_nccl():
    out = _torchrun("gpu_check_multi_gpt.py", min(_ngpu(), 8), 29532)
    assert "MULTI-GPU GPT CHECK PASS" in out


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
def test_pipeline_engine_matches_single_gpu_engine():
    n = 4 if _ngpu() >= 4 else 2
    out = _torchrun("gpu_check_pp.py", n, 29533)
    assert "PIPELINE CHECK PASS" in out

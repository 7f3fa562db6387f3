"""CPU tier: the parts of the bench / entry-point contract that do not need a GPU."""
import json
import subprocess
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parents[1]


def _run(*args, timeout=300):
    return subprocess.run([sys.executable, *args], cwd=REPO, capture_output=True, text=True, timeout=timeout)


def test_reference_arm_reports_unavailable_and_exits_zero():
    r = _run("bench.py", "--impl", "reference")
    assert r.returncode == 0, r.stderr[-500:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and "unavailable" in line and "\n" not in line["unavailable"]


def test_bench_lists_the_baseline_configurations():
    r = _run("bench.py", "--help")
    assert r.returncode == 0
    for cfg in ("mlp_dp", "mlp_fsdp", "gpt2s_dp", "gpt2s_fsdp", "gpt2m_2d", "gpt2s_pp4"):
        assert cfg in r.stdout
    for flag in ("--gpus", "--steps", "--warmup", "--impl"):
        assert flag in r.stdout


def test_graft_entry_exposes_build_and_smoke():
    import importlib
    g = importlib.import_module("__graft_entry__")
    assert callable(g.build) and callable(g.smoke)

"""CPU tier: the C ABI header, the built library and the ctypes declarations stay in sync.

The kernel library is plain CUDA behind ``include/b200dist.h``; it is cross-compiled here (no GPU
needed) and can be loaded on a CPU-only box because the CUDA runtime is linked statically.
"""
import ctypes
import re
import shutil
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parents[1]
HEADER = (REPO / "include" / "b200dist.h").read_text()


def _lib():
    from jax_distributed_tuts_b200.ops import _build, _lib
    if not _build.LIB_PATH.exists():
        if shutil.which("nvcc") is None and not Path("/usr/local/cuda/bin/nvcc").exists():
            pytest.skip("library not built and nvcc not available")
        _build.build()
    return _lib


def test_abi_version_matches_header_and_library():
    lib = _lib()
    header_version = int(re.search(r"#define\s+B200_ABI_VERSION\s+(\d+)", HEADER).group(1))
    assert header_version == lib.ABI_VERSION == lib.lib().b200_abi_version()


def test_every_exported_function_is_declared_for_ctypes():
    lib = _lib()
    handle = lib.lib()
    names = re.findall(r"^(?:int|long long|void|const char\s*\*)\s*(b200_\w+)\s*\(", HEADER, flags=re.M)
    assert len(names) >= 30, names
    missing = [n for n in names if not hasattr(handle, n)]
    assert not missing, f"declared in the header but not exported by the library: {missing}"
    undeclared = []
    for n in names:
        fn = getattr(handle, n)
        takes_args = not re.search(rf"{n}\s*\(\s*void\s*\)", HEADER)
        if takes_args and fn.argtypes is None:
            undeclared.append(n)
    assert not undeclared, f"no ctypes argtypes (arguments would be truncated to C int): {undeclared}"


def test_struct_sizes_match_the_header_layout():
    """ctypes mirrors of the ABI structs: spot-check the sizes implied by the header's field lists."""
    lib = _lib()
    assert ctypes.sizeof(lib.AdamWHParams) == 6 * 4
    assert ctypes.sizeof(lib.GatherSeg) == 7 * 8
    assert ctypes.sizeof(lib.StepEnd) == 4 * 8
    # b200_gemm_args ends with the fused cross-entropy fields (pointer, pointer, float + tail padding)
    fields = [f[0] for f in lib.GemmArgs._fields_]
    assert fields[-3:] == ["xent_labels", "xent_metrics", "xent_scale"]
    assert ctypes.sizeof(lib.GemmArgs) % 8 == 0


def test_errors_are_reported_not_swallowed():
    lib = _lib()
    handle = lib.lib()
    # an invalid problem is refused with a message before anything touches a device
    rc = handle.b200_attention_fwd(None, None, None, 1, 100, 1, None)      # seq not a multiple of 128
    assert rc != 0
    assert b"multiple of 128" in handle.b200_last_error()
    with pytest.raises(lib.B200Error):
        lib.check(rc)


def test_comm_entry_points_validate_before_touching_a_device():
    """Argument errors of the fused-collective entry points are reported by the host code (no GPU needed)."""
    lib = _lib()
    h = lib.lib()
    assert h.b200_comm_wait_published(None, 3, 2, None) != 0 and b"bad rank" in h.b200_last_error()
    assert h.b200_comm_wait_published(None, 0, 1, None) == 0            # one rank: nothing to wait for, no launch
    assert h.b200_cast_zero_wait(None, None, -1, None, 0, None, 0, None, None, 0, 1, None) != 0
    assert b"negative size" in h.b200_last_error()
    a = lib.CollectiveArgs()
    a.rank, a.world, a.mode, a.n = 0, 9, 0, 16
    assert h.b200_fused_collective(ctypes.byref(a), None) != 0 and b"world" in h.b200_last_error()
    a.world, a.mode = 2, 7
    assert h.b200_fused_collective(ctypes.byref(a), None) != 0 and b"unknown mode" in h.b200_last_error()
    a.mode, a.n = 0, 10                                                   # not a multiple of 4 x world
    assert h.b200_fused_collective(ctypes.byref(a), None) != 0 and b"multiple of 4" in h.b200_last_error()
    assert h.b200_fused_update_emulated(None, 0, None, None) != 0 and b"emulated ranks" in h.b200_last_error()
    assert int(h.b200_fused_update_emulated_scratch_bytes()) > 0
    assert ctypes.sizeof(lib.Seg) == 48 and lib.SIG_PUB == 128 * 8 + 128
    assert h.b200_source_hash().decode() not in ("", "unstamped")

"""Builds csrc/*.cu into the in-tree shared library ``_lib/libb200dist.so`` for sm_100a.

The library is plain CUDA behind a C ABI (``include/b200dist.h``): no PyTorch headers, so a
full rebuild takes well under a minute and nvcc can cross-compile it on a box without a GPU.
The built ``.so`` is git-ignored but travels with the working tree to the GPU box.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parents[1]
REPO = PKG_DIR.parent
CSRC = REPO / "csrc"
INCLUDE = REPO / "include"
LIB_DIR = PKG_DIR / "_lib"
LIB_PATH = LIB_DIR / "libb200dist.so"
OBJ_DIR = REPO / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-I", str(INCLUDE),
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; set NVCC or put it on PATH")


def _sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _deps_mtime() -> float:
    hdrs = list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))
    return max((h.stat().st_mtime for h in hdrs), default=0.0)


def source_hash() -> str:
    """Digest of everything the library is built from (csrc/*.cu, csrc/*.cuh, include/*.h). It is compiled
    into the library (``b200_source_hash()``) and kept beside it, so a library built from other sources -
    an older checkout, a header edited after the build - is rebuilt instead of loaded."""
    import hashlib
    h = hashlib.sha256()
    for f in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))):
        h.update(f.name.encode())
        h.update(f.read_bytes())
    return h.hexdigest()[:16]


HASH_PATH = LIB_DIR / "libb200dist.so.hash"


def needs_build() -> bool:
    if not LIB_PATH.exists() or not HASH_PATH.exists():
        return True
    return HASH_PATH.read_text().strip() != source_hash()


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libb200dist.so``. Returns its path."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = _nvcc()
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    hdr_m = _deps_mtime()
    digest = source_hash()

    def compile_one(src: Path) -> Path:
        obj = OBJ_DIR / (src.stem + ".o")
        stamped = src.name == "common.cu"          # carries the source digest: always recompiled (a second)
        if not force and not stamped and obj.exists() and obj.stat().st_mtime > max(src.stat().st_mtime, hdr_m):
            return obj
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
        if stamped:
            cmd.insert(1, f'-DB200_SOURCE_HASH="{digest}"')
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        objs = list(ex.map(compile_one, _sources()))
    tmp = LIB_PATH.with_suffix(".so.tmp")
    # cudart is linked statically (nvcc default): the library has no soname dependency on a
    # particular libcudart and shares the primary context with PyTorch through the driver.
    cmd = [nvcc, "-shared", "-o", str(tmp), *map(str, objs)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, LIB_PATH)
    HASH_PATH.write_text(digest + "\n")
    return LIB_PATH


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(p)

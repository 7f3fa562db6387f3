"""ctypes binding of ``libb200dist.so`` (C ABI in ``include/b200dist.h``).

The library is loaded lazily. On a machine with a GPU a missing or stale library is an error
(no silent fallback to PyTorch ops): the caller is told to run ``__graft_entry__.build()``.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional

from . import _build

ABI_VERSION = 17

_lib: Optional[C.CDLL] = None


class GemmArgs(C.Structure):
    _fields_ = [
        ("a", C.c_void_p),
        ("b_src", C.c_void_p * 8),
        ("d", C.c_void_p),
        ("m", C.c_int32), ("n", C.c_int32), ("k", C.c_int32),
        ("lda", C.c_int32), ("ldb", C.c_int32), ("ldd", C.c_int32),
        ("a_mn", C.c_int32), ("b_mn", C.c_int32),
        ("b_nsrc", C.c_int32), ("b_shard_mode", C.c_int32), ("b_shard_len", C.c_int32),
        ("b_rot", C.c_int32),
        ("d_fp32", C.c_int32),
        ("accumulate", C.c_int32),
        ("alpha", C.c_float),
        ("bias", C.c_void_p),
        ("pre", C.c_void_p),
        ("ldpre", C.c_int32),
        ("act", C.c_int32),
        ("resid", C.c_void_p),
        ("ldres", C.c_int32),
        ("bwd", C.c_int32),
        ("aux", C.c_void_p),
        ("ldaux", C.c_int32),
        ("drop_key", C.c_void_p),
        ("drop_salt", C.c_uint64),
        ("drop_rate", C.c_float),
        ("block_n", C.c_int32),
        ("drop_nsalt", C.c_int32), ("drop_rows_per_salt", C.c_int32),
        ("drop_salts", C.c_uint64 * 16),
        ("pair_mode", C.c_int32),
        ("split_k", C.c_int32),
        ("dbias", C.c_void_p),
        ("xent_labels", C.c_void_p),
        ("xent_metrics", C.c_void_p),
        ("xent_scale", C.c_float),
    ]


class AdamWHParams(C.Structure):
    _fields_ = [
        ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float),
        ("eps", C.c_float), ("weight_decay", C.c_float), ("grad_scale", C.c_float),
    ]


class StepEnd(C.Structure):
    _fields_ = [
        ("rng_key", C.c_void_p), ("salt_next", C.c_uint64),
        ("step_count", C.c_void_p), ("done_word", C.c_void_p),
    ]


class CommCtx(C.Structure):
    _fields_ = [
        ("rank", C.c_int32), ("world", C.c_int32),
        ("peer_grad", C.POINTER(C.c_void_p)),
        ("peer_param", C.POINTER(C.c_void_p)),
        ("peer_bf16", C.POINTER(C.c_void_p)),
        ("peer_signal", C.POINTER(C.c_void_p)),
        ("mc_grad", C.c_void_p),
        ("mc_param", C.c_void_p),
        ("mc_bf16", C.c_void_p),
    ]


class Seg(C.Structure):
    _fields_ = [
        ("src_off", C.c_int64), ("src_pitch", C.c_int64), ("dst_off", C.c_int64), ("loc_off", C.c_int64),
        ("rows", C.c_int32), ("row_len", C.c_int32), ("flags", C.c_int32), ("reserved", C.c_int32),
    ]


SEG_ORDERED = 1     # reduce with peer loads in rank order (segments that several ranks reduce redundantly)


class FusedUpdateArgs(C.Structure):
    _fields_ = [
        ("ctx", C.POINTER(CommCtx)), ("segs_dev", C.c_void_p),
        ("nseg", C.c_int32), ("broadcast", C.c_int32), ("publish_fp32", C.c_int32), ("nmetrics", C.c_int32),
        ("blocks", C.c_int32),
        ("p_local", C.c_void_p), ("bf16_local", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p),
        ("hp", C.POINTER(AdamWHParams)), ("step_count", C.c_void_p), ("step_metrics", C.c_void_p),
        ("metrics_acc", C.c_void_p), ("end", C.POINTER(StepEnd)),
    ]


class CollectiveArgs(C.Structure):
    _fields_ = [
        ("rank", C.c_int32), ("world", C.c_int32), ("mode", C.c_int32), ("blocks", C.c_int32),
        ("peer_buf", C.POINTER(C.c_void_p)), ("peer_signal", C.POINTER(C.c_void_p)),
        ("mc_buf", C.c_void_p), ("out", C.c_void_p), ("n", C.c_int64), ("scale", C.c_float),
    ]


class GatherSeg(C.Structure):
    _fields_ = [("src_off", C.c_int64), ("dst_off", C.c_int64), ("row_bytes", C.c_int64), ("rows", C.c_int64),
                ("src_pitch", C.c_int64), ("dst_pitch", C.c_int64), ("peer_stride", C.c_int64)]


SIGNAL_WORDS = 8192
COMM_MAX_BLOCKS = 128
# word offsets inside a signal buffer (csrc/comm.cuh)
SIG_READY, SIG_EPOCH, SIG_PUB = 0, 128 * 8, 128 * 8 + 128


class B200Error(RuntimeError):
    pass


def _declare(lib: C.CDLL) -> None:
    vp, i32, i64, f32, u64 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_uint64
    lib.b200_last_error.restype = C.c_char_p
    lib.b200_source_hash.restype = C.c_char_p
    lib.b200_source_hash.argtypes = []
    lib.b200_abi_version.restype = C.c_int
    lib.b200_launch_count.restype = C.c_longlong
    lib.b200_reset_launch_count.restype = None
    lib.b200_set_pdl.restype = None
    lib.b200_set_pdl.argtypes = [C.c_int]
    lib.b200_comm_set_timeout_ms.restype = None
    lib.b200_comm_set_timeout_ms.argtypes = [C.c_longlong]
    lib.b200_comm_set_defer_publish.restype = None
    lib.b200_comm_set_defer_publish.argtypes = [C.c_int]
    lib.b200_comm_get_defer_publish.restype = C.c_int
    lib.b200_comm_get_defer_publish.argtypes = []
    lib.b200_comm_set_stamps.restype = None
    lib.b200_comm_set_stamps.argtypes = [C.c_void_p]
    lib.b200_fused_update_emulated_scratch_bytes.restype = C.c_longlong
    lib.b200_fused_update_emulated_scratch_bytes.argtypes = []
    if hasattr(lib, "b200_attention_bwd_scratch_bytes"):
        lib.b200_attention_bwd_scratch_bytes.restype = C.c_longlong
        lib.b200_attention_bwd_scratch_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    sigs = {
        "b200_gemm_bf16": [C.POINTER(GemmArgs), vp],
        "b200_softmax_xent": [vp, i32, vp, i32, i32, f32, vp, i32, vp, vp, vp],
        "b200_softmax_xent_dbias": [vp, i32, vp, i32, i32, f32, vp, i32, vp, vp, vp, vp],
        "b200_layernorm_bwd_dxsum": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, vp],
        "b200_softmax_xent_bf16_inplace": [vp, i32, vp, i32, i32, f32, vp, vp, vp],
        "b200_colsum_bf16": [vp, i32, i32, i32, vp, i32, vp],
        "b200_adamw": [vp, vp, vp, vp, vp, i64, C.POINTER(AdamWHParams), vp, i32, i32, vp],
        "b200_cast_f32_bf16": [vp, vp, i64, vp],
        "b200_cast_bf16_f32": [vp, vp, i64, vp],
        "b200_scale_f32": [vp, i64, f32, vp],
        "b200_rng_split": [vp, vp, u64, u64, vp],
        "b200_rng_split_bump": [vp, vp, u64, u64, vp, vp],
        "b200_cast_zero": [vp, vp, i64, vp, i64, vp, i64, vp],
        "b200_cast_zero_wait": [vp, vp, i64, vp, i64, vp, i64, vp, C.POINTER(C.c_void_p), i32, i32, vp],
        "b200_fused_update": [C.POINTER(FusedUpdateArgs), vp],
        "b200_fused_update_emulated": [C.POINTER(FusedUpdateArgs), i32, vp, vp],
        "b200_fused_collective": [C.POINTER(CollectiveArgs), vp],
        "b200_fused_collective_emulated": [C.POINTER(CollectiveArgs), i32, vp, vp],
        "b200_comm_wait_published": [vp, i32, i32, vp],
        "b200_act_dropout_fwd": [vp, vp, i64, i32, i32, i32, vp, u64, f32, vp],
        "b200_act_dropout_bwd": [vp, vp, vp, i64, i32, i32, i32, vp, u64, f32, vp],
        "b200_layernorm_fwd": [vp, vp, vp, vp, vp, vp, i32, i32, f32, vp],
        "b200_layernorm_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, vp],
        "b200_embedding_fwd": [vp, vp, vp, vp, i32, i32, i32, vp],
        "b200_embedding_bwd": [vp, vp, vp, vp, i32, i32, i32, vp],
        "b200_embedding_fwd_sharded": [vp, C.POINTER(C.c_void_p), i32, i64, C.POINTER(C.c_void_p), i32, i64, vp,
                                       i32, i32, i32, vp],
        "b200_attention_fwd": [vp, vp, vp, i32, i32, i32, vp],
        "b200_attention_bwd": [vp, vp, vp, vp, vp, vp, i32, i32, i32, vp],
        "b200_attention_bwd_noalloc": [vp, vp, vp, vp, vp, i32, i32, i32, vp],
        "b200_fused_reduce_adamw": [C.POINTER(CommCtx), vp, i32, i32, vp, vp, vp, vp,
                                    C.POINTER(AdamWHParams), vp, i32, vp, vp, i32, i32, vp],
        "b200_fused_reduce_adamw_end": [C.POINTER(CommCtx), vp, i32, i32, vp, vp, vp, vp,
                                        C.POINTER(AdamWHParams), vp, i32, vp, vp, i32, i32, C.POINTER(StepEnd), vp],
        "b200_signal_barrier": [C.POINTER(CommCtx), vp],
        "b200_peer_gather": [C.POINTER(C.c_void_p), i32, vp, i32, vp, i32, vp],
    }
    for name, argtypes in sigs.items():
        fn = getattr(lib, name, None)
        if fn is None:  # entry point not built yet
            continue
        fn.argtypes = argtypes
        fn.restype = C.c_int


def lib() -> C.CDLL:
    """Returns the loaded kernel library, building it first if the sources are newer."""
    global _lib
    if _lib is not None:
        return _lib
    path = Path(os.environ["B200DIST_LIB"]) if os.environ.get("B200DIST_LIB") else _build.LIB_PATH   # override: A/B builds
    if path == _build.LIB_PATH and _build.needs_build():
        # missing, or built from other sources than this tree's (another checkout, a header edited since):
        # rebuild where a compiler exists, otherwise fall through to the checks below
        try:
            _build.build()
        except Exception:  # noqa: BLE001
            pass
    if not path.exists():
        raise B200Error(
            f"{path} is missing: build the CUDA library first "
            "(python -c 'import __graft_entry__ as g; g.build()')"
        )
    handle = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
    _declare(handle)
    got = handle.b200_abi_version()
    if got != ABI_VERSION:
        raise B200Error(f"{path} has ABI version {got}, expected {ABI_VERSION}: rebuild it")
    if path == _build.LIB_PATH and hasattr(handle, "b200_source_hash"):
        handle.b200_source_hash.restype = C.c_char_p
        built_from = handle.b200_source_hash().decode()
        if built_from != _build.source_hash():
            raise B200Error(f"{path} was built from other sources (digest {built_from}, this tree {_build.source_hash()}): "
                            "rebuild it (python -c 'import __graft_entry__ as g; g.build()')")
    _lib = handle
    return handle


def available() -> bool:
    try:
        lib()
        return True
    except (OSError, B200Error):
        return False


def check(rc: int) -> None:
    if rc != 0:
        msg = lib().b200_last_error()
        raise B200Error(msg.decode("utf-8", "replace") if msg else f"b200 error {rc}")


def launch_count() -> int:
    return int(lib().b200_launch_count())


def reset_launch_count() -> None:
    lib().b200_reset_launch_count()


def set_pdl(mode: int) -> None:
    """1 / 0 forces programmatic dependent launch on / off for later launches, -1 restores the default."""
    lib().b200_set_pdl(int(mode))

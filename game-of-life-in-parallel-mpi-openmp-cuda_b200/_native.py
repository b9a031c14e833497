"""ctypes binding of libgolb200.so (the C ABI in include/gol_b200.h).

The library is built in-tree by ``make lib`` (see ``__graft_entry__.build``); it is loaded from
the repository root so that the driver sees the in-tree ``.so`` in the process map.  Loading
works on a machine without a GPU (the CUDA runtime is linked statically and only touched by
the device entry points), which is what the CPU-only tests rely on.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

REPO_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("GOL_LIB") or os.path.join(REPO_ROOT, "libgolb200.so")  # GOL_LIB: kernel-variant experiments

GHOST_ROWS = 16
GHOST_WORDS32 = 4
MIN_FAST_WIDTH = 128
MIN_FAST_HEIGHT = 16
EXIT_LIMIT, EXIT_EMPTY, EXIT_SIMILAR = 0, 1, 2
EXIT_NAMES = {0: "limit", 1: "empty", 2: "similar"}
FLAG_FORCE_GENERIC = 1


class GolError(RuntimeError):
    """Raised when a native call fails; carries gol_last_error()."""


class Layout(C.Structure):
    _fields_ = [("width", C.c_int64), ("height", C.c_int64), ("ghost_rows", C.c_int32),
                ("nq", C.c_int32), ("nqf", C.c_int32), ("rem", C.c_int32), ("nw32", C.c_int32),
                ("pitch32", C.c_int32), ("rows_padded", C.c_int64), ("buffer_words32", C.c_int64)]


class Result(C.Structure):
    _fields_ = [("generations", C.c_int32), ("exit_reason", C.c_int32),
                ("generations_computed", C.c_int64), ("loop_ms", C.c_double),
                ("read_ms", C.c_double), ("write_ms", C.c_double), ("population", C.c_uint64)]


class Config(C.Structure):
    _fields_ = [("width", C.c_int64), ("height", C.c_int64), ("grid_rows", C.c_int32),
                ("grid_cols", C.c_int32), ("proc_rank", C.c_int32), ("proc_count", C.c_int32),
                ("device", C.c_int32), ("gen_limit", C.c_int32), ("similarity_frequency", C.c_int32),
                ("k_max", C.c_int32), ("batch_gens", C.c_int32), ("flags", C.c_uint32),
                ("start_generation", C.c_int32), ("reserved", C.c_int32)]


_lock = threading.Lock()
_lib = None


def build(force: bool = False) -> str:
    """Compile libgolb200.so (and the CLI binaries) for sm_100a with the in-tree Makefile."""
    if force or not os.path.exists(LIB_PATH):
        subprocess.run(["make", "-C", REPO_ROOT, "-j8", "all"], check=True,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    return LIB_PATH


def _proto(lib):
    p64 = C.POINTER(C.c_uint64)
    vp, i64, i32, u64 = C.c_void_p, C.c_int64, C.c_int, C.c_uint64
    sigs = {
        "gol_last_error": (C.c_char_p, []),
        "gol_version": (i32, []),
        "gol_cuda_available": (i32, []),
        "gol_layout_init": (i32, [C.POINTER(Layout), i64, i64]),
        "gol_cpu_step": (i32, [vp, i64, i64, i64, i32]),
        "gol_cpu_run": (i32, [vp, i64, i64, i32, i32, i32, C.POINTER(Result)]),
        "gol_text_parse": (i32, [C.c_char_p, C.c_size_t, i64, i64, i32, vp]),
        "gol_text_format": (i32, [vp, i64, i64, vp]),
        "gol_file_read_packed": (i32, [C.c_char_p, i64, i64, i32, vp]),
        "gol_file_write_packed": (i32, [C.c_char_p, vp, i64, i64]),
        "gol_random_packed": (i32, [vp, i64, i64, u64]),
        "gol_random_text_file": (i32, [C.c_char_p, i64, i64, u64]),
        "gol_write_pbm": (i32, [C.c_char_p, vp, i64, i64]),
        "gol_op_step_dense": (i32, [vp, vp, i64, i64, vp, vp]),
        "gol_op_flags_dense": (i32, [vp, vp, i64, i64, vp, vp]),
        "gol_op_pack_cells": (i32, [vp, vp, i64, i64, vp]),
        "gol_op_unpack_cells": (i32, [vp, vp, i64, i64, vp]),
        "gol_op_parse_text": (i32, [vp, vp, i64, i64, i64, vp, vp]),
        "gol_op_parse_text_n": (i32, [vp, i64, vp, i64, i64, i64, vp, vp]),
        "gol_op_format_text": (i32, [vp, vp, i64, i64, i64, i32, vp]),
        "gol_op_random_dense": (i32, [vp, i64, i64, i64, i64, u64, vp]),
        "gol_op_popcount_dense": (i32, [vp, i64, i64, vp, vp]),
        "gol_config_default": (None, [C.POINTER(Config), i64, i64]),
        "gol_engine_create": (i32, [C.POINTER(Config), C.POINTER(vp)]),
        "gol_engine_destroy": (None, [vp]),
        "gol_engine_set_stream": (i32, [vp, vp]),
        "gol_engine_num_domains": (i32, [vp]),
        "gol_engine_num_local_domains": (i32, [vp]),
        "gol_engine_local_domain_id": (i32, [vp, i32]),
        "gol_engine_domain_extent": (i32, [vp, i32, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]),
        "gol_engine_domain_neighbors": (i32, [vp, i32, C.POINTER(C.c_int32 * 8)]),
        "gol_engine_uses_fast_path": (i32, [vp]),
        "gol_engine_ipc_blob_size": (C.c_size_t, [vp]),
        "gol_engine_ipc_export": (i32, [vp, vp]),
        "gol_engine_ipc_import": (i32, [vp, vp]),
        "gol_engine_load_packed": (i32, [vp, vp]),
        "gol_engine_load_packed_device": (i32, [vp, vp, i64, i64, i64, i64]),
        "gol_engine_stage_packed": (i32, [vp, vp]),
        "gol_engine_commit_staged": (i32, [vp]),
        "gol_engine_randomize": (i32, [vp, u64]),
        "gol_engine_load_text_file": (i32, [vp, C.c_char_p, i32]),
        "gol_engine_store_text_file": (i32, [vp, C.c_char_p]),
        "gol_engine_download_packed": (i32, [vp, vp]),
        "gol_engine_prepare": (i32, [vp]),
        "gol_engine_step": (i32, [vp, i64]),
        "gol_engine_run": (i32, [vp, C.POINTER(Result)]),
        "gol_engine_sync": (i32, [vp]),
        "gol_engine_launch_count": (i64, [vp]),
        "gol_engine_chunk_rows": (i32, [vp]),
        "gol_engine_stage_domain": (i32, [vp, i32, vp]),
        "gol_engine_download_domain": (i32, [vp, i32, vp]),
        "gol_engine_hash": (i32, [vp, p64]),
        "gol_hash_packed": (i32, [vp, i64, i64, p64]),
        "gol_engine_uses_deep_halo": (i32, [vp]),
        "gol_engine_trace_dump": (i32, [vp, C.c_char_p]),
        "gol_engine_population_static": (i32, [vp, p64]),
        "gol_engine_population": (i32, [vp, p64]),
        "gol_engine_domain_buffer": (i32, [vp, i32, C.POINTER(vp), C.POINTER(Layout)]),
        "gol_engine_last_ms": (C.c_double, [vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


def lib():
    """The loaded native library (built on first use if missing)."""
    global _lib
    with _lock:
        if _lib is None:
            build()
            _lib = _proto(C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL))
        return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().gol_last_error().decode("utf-8", "replace")
        raise GolError(f"{what}: {msg}" if what else msg)


def cuda_available() -> bool:
    return bool(lib().gol_cuda_available())


def require_cuda() -> None:
    """Fail loudly: the device ops have no eager / PyTorch fallback."""
    if not cuda_available():
        raise GolError("libgolb200: no usable CUDA device; the device ops have no CPU fallback "
                       "(use gol_b200.ops.cpu for the CPU oracle)")

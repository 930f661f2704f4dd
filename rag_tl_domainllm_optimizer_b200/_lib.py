"""ctypes binding of ``libb2r.so`` (the C ABI declared in ``include/b2r.h``).

There is no fallback: if the CUDA library has not been built, importing the binding raises, and
if no CUDA device is present every compute call fails with the library's own error.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb2r.so")

B2R_OK = 0
ERR_NAMES = {-1: "B2R_ERR_ARG", -2: "B2R_ERR_CUDA", -3: "B2R_ERR_OOM", -4: "B2R_ERR_STATE", -5: "B2R_ERR_LIMIT"}

METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1
STORE_F32 = 0
STORE_BF16 = 1
ENGINE_AUTO, ENGINE_FMA, ENGINE_UMMA, ENGINE_UMMA3, ENGINE_SHADOW = 0, 1, 2, 3, 4

# every symbol include/b2r.h declares (tests check that the library exports all of them)
EXPORTS = [
    "b2r_last_error", "b2r_version", "b2r_device_count", "b2r_index_create", "b2r_index_free", "b2r_index_add",
    "b2r_index_reserve", "b2r_index_reset", "b2r_index_ntotal", "b2r_index_reconstruct", "b2r_index_search",
    "b2r_index_search_ex", "b2r_index_search_packed", "b2r_merge_packed", "b2r_index_set_engine",
    "b2r_index_set_pad", "b2r_index_last_stats", "b2r_index_last_score_ms", "b2r_index_score_time", "b2r_normalize_l2", "b2r_cosine_rows",
    "b2r_cosine_rows_noalloc", "b2r_index_export_rows", "b2r_index_import_rows", "b2r_index_set_shadow", "b2r_index_set_verify",
    "b2r_merge_packed_strided", "b2r_nccl_unique_id", "b2r_index_set_comm", "b2r_index_comm_init", "b2r_index_search_sharded",
]


class SearchStats(ctypes.Structure):
    _fields_ = [("nq", ctypes.c_int64), ("k", ctypes.c_int64), ("kprime", ctypes.c_int64),
                ("unverified", ctypes.c_int64), ("candidates", ctypes.c_int64), ("engine", ctypes.c_int32),
                ("launches", ctypes.c_int32), ("fallback", ctypes.c_int32), ("reserved", ctypes.c_int32),
                ("main_rows", ctypes.c_int64), ("fallback_queries", ctypes.c_int64),
                ("exhaustive_queries", ctypes.c_int64)]


class B2RError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"{ERR_NAMES.get(code, code)}: {msg}")
        self.code = code


def build(verbose: bool = False) -> str:
    """Compile csrc/ -> libb2r.so for sm_100a with nvcc (in-tree; cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc")]
    if not verbose:
        cmd.append("-s")
    subprocess.check_call(cmd)
    return LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU or PyTorch fallback for this library.")
    L = ctypes.CDLL(LIB_PATH)
    vp, i64, i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    L.b2r_last_error.restype = ctypes.c_char_p
    L.b2r_version.restype = ctypes.c_char_p
    L.b2r_device_count.restype = i32
    L.b2r_index_create.argtypes = [ctypes.POINTER(vp), i64, i32, i32, i32]
    L.b2r_index_free.argtypes = [vp]
    L.b2r_index_free.restype = None
    L.b2r_index_add.argtypes = [vp, i64, vp, i32, vp]
    L.b2r_index_reserve.argtypes = [vp, i64]
    L.b2r_index_reset.argtypes = [vp]
    L.b2r_index_ntotal.argtypes = [vp, ctypes.POINTER(i64)]
    L.b2r_index_reconstruct.argtypes = [vp, i64, vp, vp]
    L.b2r_index_export_rows.argtypes = [vp, i64, i64, vp, vp]
    L.b2r_index_import_rows.argtypes = [vp, i64, vp, vp]
    L.b2r_index_search.argtypes = [vp, i64, vp, i64, vp, vp, vp]
    L.b2r_index_search_ex.argtypes = [vp, i64, vp, i64, i32, i64, vp, vp, vp]
    L.b2r_index_search_packed.argtypes = [vp, i64, vp, i64, i32, i64, vp, vp]
    L.b2r_merge_packed.argtypes = [i32, i64, i64, i64, vp, vp, vp, i32, vp]
    L.b2r_merge_packed_strided.argtypes = [i32, i64, i64, i64, vp, i64, vp, vp, i32, vp]
    L.b2r_nccl_unique_id.argtypes = [vp]
    L.b2r_index_set_comm.argtypes = [vp, vp, i32, i32, i64]
    L.b2r_index_comm_init.argtypes = [vp, vp, i32, i32, i64]
    L.b2r_index_search_sharded.argtypes = [vp, i64, vp, i64, i32, vp, vp, vp]
    L.b2r_index_set_engine.argtypes = [vp, i32]
    L.b2r_index_set_pad.argtypes = [vp, i32]
    L.b2r_index_set_shadow.argtypes = [vp, i32]
    L.b2r_index_set_verify.argtypes = [vp, i32]
    L.b2r_index_last_stats.argtypes = [vp, ctypes.POINTER(SearchStats)]
    L.b2r_index_last_score_ms.argtypes = [vp, ctypes.POINTER(ctypes.c_float)]
    L.b2r_index_score_time.argtypes = [vp, i32, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(i64)]
    L.b2r_normalize_l2.argtypes = [i64, i64, vp, i32, vp]
    L.b2r_cosine_rows.argtypes = [i64, i64, i64, vp, vp, vp, vp, i32, vp]
    L.b2r_cosine_rows_noalloc.argtypes = [i64, i64, i64, vp, vp, vp, vp, i32, vp]
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != B2R_OK:
        raise B2RError(rc, lib().b2r_last_error().decode("utf-8", "replace"))

"""CPU tests of the C-ABI boundary: the library loads, exports everything include/b2r.h declares, and
refuses to compute without a CUDA device (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

from rag_tl_domainllm_optimizer_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(ROOT, "include", "b2r.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(b2r_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    L = _lib.lib()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/b2r.h but not exported by libb2r.so"
    assert sorted(n for n in _lib.EXPORTS) == sorted(n for n in names if n in _lib.EXPORTS)
    missing_in_binding = [n for n in names if n not in _lib.EXPORTS]
    assert not missing_in_binding, missing_in_binding


def test_version_and_error_string():
    L = _lib.lib()
    assert L.b2r_version().startswith(b"b2r")
    assert isinstance(L.b2r_last_error(), bytes)


def test_no_cpu_fallback_without_device():
    L = _lib.lib()
    if L.b2r_device_count() > 0:
        pytest.skip("a CUDA device is present")
    h = ctypes.c_void_p()
    rc = L.b2r_index_create(ctypes.byref(h), 64, 0, 0, 0)
    assert rc == -2 and not h.value                       # B2R_ERR_CUDA, no handle
    assert b"no CUDA device" in L.b2r_last_error()
    x = np.ones((2, 8), dtype=np.float32)
    assert L.b2r_normalize_l2(2, 8, x.ctypes.data, 0, None) == -2
    assert np.all(x == 1.0)                               # untouched: nothing was computed on the CPU


def test_argument_validation_needs_no_device():
    L = _lib.lib()
    h = ctypes.c_void_p()
    assert L.b2r_index_create(None, 64, 0, 0, 0) == -1
    assert L.b2r_index_create(ctypes.byref(h), 0, 0, 0, 0) == -1
    assert L.b2r_index_create(ctypes.byref(h), 64, 7, 0, 0) == -1
    assert L.b2r_index_create(ctypes.byref(h), 64, 0, 9, 0) == -1
    assert L.b2r_index_search(None, 1, None, 1, None, None, None) == -1
    assert L.b2r_index_ntotal(None, None) == -1
    assert L.b2r_merge_packed(0, 0, 1, 1, None, None, None, 0, None) == -1
    assert L.b2r_merge_packed(0, 2, 1, 5000, None, None, None, 0, None) == -5
    assert L.b2r_normalize_l2(-1, 8, None, 0, None) == -1
    assert L.b2r_merge_packed(0, 2000, 1, 1, None, None, None, 0, None) == -5        # more lists than the merge takes
    # row-sharded entry points: argument errors come before any device or NCCL work
    assert L.b2r_index_search_sharded(None, 1, None, 1, 0, None, None, None) == -1
    assert L.b2r_index_set_comm(None, None, 0, 1, 0) == -1
    assert L.b2r_index_comm_init(None, None, 0, 1, 0) == -1
    assert L.b2r_nccl_unique_id(None) == -1


def test_python_host_raises_not_falls_back():
    import rag_tl_domainllm_optimizer_b200 as b2r
    if _lib.lib().b2r_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(b2r.B2RError):
        b2r.IndexFlatIP(32)
    with pytest.raises(b2r.ShapeError):
        b2r.IndexFlatIP(0)

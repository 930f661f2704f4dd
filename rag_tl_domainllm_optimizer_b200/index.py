"""FAISS-shaped flat indexes over the B200 C-ABI library.

This is the host-side mirror of the interface the reference's retrieval slot expects: the
``retrieved_docs`` consumed at reinforcement_learning_optimization_after_rag.py:286-288 / :405
comes, per the reference's tech stack (README.md:27-28), from a FAISS flat index behind
LangChain / LlamaIndex, which call ``index.add(x)`` and ``D, I = index.search(x, k)``.  Names,
argument meaning, shapes, dtypes, padding (-1 / -+FLT_MAX) and best-first ordering follow
``faiss.IndexFlatIP`` / ``faiss.IndexFlatL2`` / ``faiss.normalize_L2``.

All arithmetic happens in ``libb2r.so`` (hand-written sm_100a kernels).  NumPy arrays go through
the C ABI as host pointers (the library stages them); CUDA torch tensors go through as device
pointers on torch's current stream.  Nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes
from typing import Tuple

import numpy as np

from . import _lib
from ._lib import (ENGINE_AUTO, ENGINE_FMA, ENGINE_UMMA, ENGINE_UMMA3, METRIC_INNER_PRODUCT, METRIC_L2, STORE_BF16, STORE_F32,
                   B2RError, SearchStats, check)


class ShapeError(ValueError, AssertionError):
    """Bad shape / dtype / k.  Subclasses AssertionError too: faiss's Python layer asserts."""


def _is_torch(x) -> bool:
    return type(x).__module__.split(".")[0] == "torch"


def _torch():
    import torch
    return torch


def _as_f32_2d(x, d: int, what: str):
    """-> (kind, obj, ptr, n, stream, device) with obj kept alive by the caller."""
    if _is_torch(x):
        torch = _torch()
        if x.dim() != 2 or x.shape[1] != d:
            raise ShapeError(f"{what} must have shape [n, {d}], got {tuple(x.shape)}")
        if x.dtype != torch.float32:
            raise ShapeError(f"{what} must be float32, got {x.dtype}")
        x = x.contiguous()
        if x.is_cuda:
            return "cuda", x, x.data_ptr(), x.shape[0], torch.cuda.current_stream(x.device).cuda_stream, x.device.index
        return "cpu_torch", x, x.data_ptr(), x.shape[0], 0, None
    a = np.asarray(x)
    if a.ndim != 2 or a.shape[1] != d:
        raise ShapeError(f"{what} must have shape [n, {d}], got {a.shape}")
    if a.dtype != np.float32:
        raise ShapeError(f"{what} must be float32, got {a.dtype}")
    a = np.ascontiguousarray(a)
    return "numpy", a, a.ctypes.data, a.shape[0], 0, None


class IndexFlat:
    """Exact brute-force index.  ``storage='f32'`` is the exact mode, ``'bf16'`` the fast mode."""

    def __init__(self, d: int, metric: int = METRIC_L2, *, storage: str = "f32", device: int | None = None,
                 normalize: bool = False, shadow: bool = False):
        if int(d) <= 0:
            raise ShapeError(f"d must be positive, got {d}")
        if storage not in ("f32", "bf16"):
            raise ValueError("storage must be 'f32' or 'bf16'")
        if device is None:
            device = 0
            try:
                torch = _torch()
                if torch.cuda.is_available():
                    device = torch.cuda.current_device()
            except ImportError:
                pass
        self.d = int(d)
        self.metric_type = int(metric)
        self.storage = storage
        self.device = int(device)
        self.normalize = bool(normalize)   # cosine index: rows and queries are L2-normalised on the way in
        self.is_trained = True
        self.engine = ENGINE_AUTO
        self._h = ctypes.c_void_p()
        L = _lib.lib()
        check(L.b2r_index_create(ctypes.byref(self._h), self.d, self.metric_type,
                                 STORE_BF16 if storage == "bf16" else STORE_F32, self.device))
        self.shadow = bool(shadow)       # exact mode: score a bf16 copy of the rows, re-score the fp32 rows
        if self.shadow:
            check(L.b2r_index_set_shadow(self._h, 1))

    # -- life cycle -------------------------------------------------------------------------
    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().b2r_index_free(h)
            except Exception:
                pass

    @property
    def ntotal(self) -> int:
        out = ctypes.c_int64()
        check(_lib.lib().b2r_index_ntotal(self._h, ctypes.byref(out)))
        return out.value

    def reserve(self, n: int) -> None:
        check(_lib.lib().b2r_index_reserve(self._h, int(n)))

    def reset(self) -> None:
        check(_lib.lib().b2r_index_reset(self._h))

    # -- faiss API --------------------------------------------------------------------------
    def add(self, x) -> None:
        kind, obj, ptr, n, stream, dev = _as_f32_2d(x, self.d, "x")
        if kind == "cuda" and dev != self.device:
            raise ShapeError(f"x is on cuda:{dev}, the index on cuda:{self.device}")
        check(_lib.lib().b2r_index_add(self._h, n, ptr, int(self.normalize), stream))

    def search(self, x, k: int) -> Tuple[object, object]:
        k = int(k)
        if k <= 0:
            raise ShapeError(f"k must be positive, got {k}")
        kind, obj, ptr, n, stream, dev = _as_f32_2d(x, self.d, "x")
        L = _lib.lib()
        if kind == "cuda":
            torch = _torch()
            if dev != self.device:
                raise ShapeError(f"x is on cuda:{dev}, the index on cuda:{self.device}")
            D = torch.empty((n, k), dtype=torch.float32, device=obj.device)
            I = torch.empty((n, k), dtype=torch.int64, device=obj.device)
            check(L.b2r_index_search_ex(self._h, n, ptr, k, int(self.normalize), 0, D.data_ptr(), I.data_ptr(), stream))
            return D, I
        D = np.empty((n, k), dtype=np.float32)
        I = np.empty((n, k), dtype=np.int64)
        check(L.b2r_index_search_ex(self._h, n, ptr, k, int(self.normalize), 0, D.ctypes.data, I.ctypes.data, 0))
        if kind == "cpu_torch":
            torch = _torch()
            return torch.from_numpy(D), torch.from_numpy(I)
        return D, I

    def reconstruct(self, i: int) -> np.ndarray:
        out = np.empty(self.d, dtype=np.float32)
        check(_lib.lib().b2r_index_reconstruct(self._h, int(i), out.ctypes.data, 0))
        return out

    # -- shard-local search for the multi-GPU path -----------------------------------------
    def search_packed(self, x, k: int, row_offset: int = 0):
        """Shard-local top-k left packed (int64 view of uint64 keys, CUDA tensor [n,k])."""
        torch = _torch()
        kind, obj, ptr, n, stream, dev = _as_f32_2d(x, self.d, "x")
        device = torch.device("cuda", self.device)
        if kind != "cuda":
            stream = torch.cuda.current_stream(device).cuda_stream
        packed = torch.empty((n, int(k)), dtype=torch.int64, device=device)
        check(_lib.lib().b2r_index_search_packed(self._h, n, ptr, int(k), int(self.normalize), int(row_offset),
                                                 packed.data_ptr(), stream))
        return packed

    # -- row-sharded search through the library's own collective ---------------------------------
    def comm_init(self, unique_id: bytes, rank: int, world: int, row_offset: int) -> None:
        """Join the NCCL communicator of a row-sharded corpus (``unique_id`` = rank 0's ``nccl_unique_id()``)."""
        buf = ctypes.create_string_buffer(bytes(unique_id), 128) if world > 1 else None
        check(_lib.lib().b2r_index_comm_init(self._h, buf, int(rank), int(world), int(row_offset)))

    def search_sharded(self, x, k: int) -> Tuple[object, object]:
        """Collective: every rank calls it with the same queries; returns the global top-k on every rank."""
        k = int(k)
        if k <= 0:
            raise ShapeError(f"k must be positive, got {k}")
        kind, obj, ptr, n, stream, dev = _as_f32_2d(x, self.d, "x")
        L = _lib.lib()
        if kind == "cuda":
            torch = _torch()
            if dev != self.device:
                raise ShapeError(f"x is on cuda:{dev}, the index on cuda:{self.device}")
            D = torch.empty((n, k), dtype=torch.float32, device=obj.device)
            I = torch.empty((n, k), dtype=torch.int64, device=obj.device)
            check(L.b2r_index_search_sharded(self._h, n, ptr, k, int(self.normalize), D.data_ptr(), I.data_ptr(), stream))
            return D, I
        D = np.empty((n, k), dtype=np.float32)
        I = np.empty((n, k), dtype=np.int64)
        check(L.b2r_index_search_sharded(self._h, n, ptr, k, int(self.normalize), D.ctypes.data, I.ctypes.data, 0))
        if kind == "cpu_torch":
            torch = _torch()
            return torch.from_numpy(D), torch.from_numpy(I)
        return D, I

    # -- tuning / introspection ------------------------------------------------------------
    def set_engine(self, engine: int) -> None:
        check(_lib.lib().b2r_index_set_engine(self._h, int(engine)))
        self.engine = int(engine)

    def set_verify(self, on: bool) -> None:
        """Device-side re-search of unproven queries after each search (default on; off only counts them)."""
        check(_lib.lib().b2r_index_set_verify(self._h, int(bool(on))))

    def set_pad(self, pad: int) -> None:
        check(_lib.lib().b2r_index_set_pad(self._h, int(pad)))

    def last_stats(self) -> dict:
        st = SearchStats()
        check(_lib.lib().b2r_index_last_stats(self._h, ctypes.byref(st)))
        return {f: getattr(st, f) for f, _ in SearchStats._fields_}

    def last_score_ms(self) -> float:
        ms = ctypes.c_float()
        check(_lib.lib().b2r_index_last_score_ms(self._h, ctypes.byref(ms)))
        return ms.value


    def score_time(self, reset: bool = True):
        """(total ms, searches) spent in the scoring kernel(s) since the last reset (<= 256 most recent)."""
        tot, cnt = ctypes.c_double(), ctypes.c_int64()
        check(_lib.lib().b2r_index_score_time(self._h, int(reset), ctypes.byref(tot), ctypes.byref(cnt)))
        return tot.value, cnt.value


class IndexFlatIP(IndexFlat):
    def __init__(self, d: int, **kw):
        super().__init__(d, METRIC_INNER_PRODUCT, **kw)


class IndexFlatL2(IndexFlat):
    def __init__(self, d: int, **kw):
        super().__init__(d, METRIC_L2, **kw)


def merge_packed(packed, metric: int = METRIC_INNER_PRODUCT):
    """[G, n, k] packed keys (CUDA int64 tensor) -> (D [n,k] float32, I [n,k] int64)."""
    torch = _torch()
    if packed.dim() != 3 or not packed.is_cuda or packed.dtype != torch.int64:
        raise ShapeError("packed must be a CUDA int64 tensor of shape [G, n, k]")
    packed = packed.contiguous()
    G, n, k = packed.shape
    D = torch.empty((n, k), dtype=torch.float32, device=packed.device)
    I = torch.empty((n, k), dtype=torch.int64, device=packed.device)
    check(_lib.lib().b2r_merge_packed(int(metric), G, n, k, packed.data_ptr(), D.data_ptr(), I.data_ptr(),
                                      packed.device.index, torch.cuda.current_stream(packed.device).cuda_stream))
    return D, I


def nccl_unique_id() -> bytes:
    """A fresh 128-byte ncclUniqueId (rank 0 calls this and sends it to the other ranks)."""
    buf = ctypes.create_string_buffer(128)
    check(_lib.lib().b2r_nccl_unique_id(buf))
    return buf.raw


def merge_packed_strided(gathered, n: int, k: int, metric: int = METRIC_INNER_PRODUCT):
    """[G, >= n*k] payloads whose first n*k words are packed hits (CUDA int64) -> (D [n,k], I [n,k])."""
    torch = _torch()
    if gathered.dim() != 2 or not gathered.is_cuda or gathered.dtype != torch.int64 or gathered.shape[1] < n * k:
        raise ShapeError("gathered must be a CUDA int64 tensor of shape [G, >= n*k]")
    gathered = gathered.contiguous()
    D = torch.empty((n, k), dtype=torch.float32, device=gathered.device)
    I = torch.empty((n, k), dtype=torch.int64, device=gathered.device)
    check(_lib.lib().b2r_merge_packed_strided(int(metric), gathered.shape[0], n, k, gathered.data_ptr(),
                                              gathered.shape[1], D.data_ptr(), I.data_ptr(), gathered.device.index,
                                              torch.cuda.current_stream(gathered.device).cuda_stream))
    return D, I


def normalize_L2(x, device: int = 0) -> None:
    """In-place row normalisation (faiss.normalize_L2); zero rows stay zero."""
    if _is_torch(x):
        torch = _torch()
        if x.dim() != 2 or x.dtype != torch.float32 or not x.is_contiguous():
            raise ShapeError("x must be a contiguous float32 matrix")
        stream = torch.cuda.current_stream(x.device).cuda_stream if x.is_cuda else 0
        dev = x.device.index if x.is_cuda else device
        check(_lib.lib().b2r_normalize_l2(x.shape[0], x.shape[1], x.data_ptr(), dev, stream))
        return
    if not isinstance(x, np.ndarray) or x.ndim != 2 or x.dtype != np.float32 or not x.flags.c_contiguous:
        raise ShapeError("x must be a C-contiguous float32 matrix")
    check(_lib.lib().b2r_normalize_l2(x.shape[0], x.shape[1], x.ctypes.data, device, 0))


def cosine_rows(resp, docs, device: int = 0):
    """Batched reward cosine (reference :63-79): resp [n,d], docs [n,m,d] -> (cos [n,m], max [n])."""
    resp = np.ascontiguousarray(resp, dtype=np.float32)
    docs = np.ascontiguousarray(docs, dtype=np.float32)
    n, d = resp.shape
    m = docs.shape[1] if docs.ndim == 3 and docs.size else 0
    cos = np.empty((n, m), dtype=np.float32)
    mx = np.empty(n, dtype=np.float32)
    check(_lib.lib().b2r_cosine_rows(n, m, d, resp.ctypes.data, docs.ctypes.data if m else None,
                                     cos.ctypes.data if m else None, mx.ctypes.data, device, 0))
    return cos, mx

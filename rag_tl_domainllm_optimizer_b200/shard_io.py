"""On-disk shard format of a flat index and a faiss flat-index file reader/writer.

``save_shard`` / ``load_shard`` write one rank's rows in their STORED representation (fp32, or bf16 in fast
mode), so a loaded index scores bit-identically to the one that was saved; one file per rank of a row-sharded
corpus, with the rank's ``row_offset`` in the header.

Shard file, little endian::

    0   8s  magic  b"B2RSHRD1"
    8   q   d
    16  q   ntotal
    24  q   row_offset      global id of the first row
    32  i   metric          0 inner product, 1 L2
    36  i   storage         0 float32, 1 bfloat16
    40  i   normalized      1 if rows were L2-normalised on ingest
    44  I   crc32           zlib.crc32 of the payload
    48  16x reserved
    64  payload             ntotal * d elements, row-major, dense

``write_faiss_flat`` / ``read_faiss_flat`` use the layout of faiss ``write_index`` for ``IndexFlat`` (fourcc
``IxFI`` / ``IxF2``, d, ntotal, two dummies, is_trained, metric_type, then the float vector with its element
count).  That layout is restated from faiss's published ``index_write.cpp``; faiss is not installable in the
build environment, so the two functions are only tested against each other — check them against a real faiss
file before relying on them for migration.
"""
from __future__ import annotations

import struct
import zlib

import numpy as np

from . import _lib
from ._lib import METRIC_INNER_PRODUCT, METRIC_L2, check
from .index import IndexFlat

MAGIC = b"B2RSHRD1"
_HDR = struct.Struct("<8sqqqiiiI16x")
_CHUNK_BYTES = 256 << 20


def _export(index: IndexFlat, row0: int, n: int) -> np.ndarray:
    dt = np.float32 if index.storage == "f32" else np.uint16
    out = np.empty((n, index.d), dtype=dt)
    check(_lib.lib().b2r_index_export_rows(index._h, row0, n, out.ctypes.data, 0))
    return out


def save_shard(index: IndexFlat, path: str, row_offset: int = 0) -> None:
    n, d = index.ntotal, index.d
    esz = 4 if index.storage == "f32" else 2
    rows_per = max(1, _CHUNK_BYTES // (d * esz))
    crc = 0
    with open(path, "wb") as f:
        f.write(b"\0" * _HDR.size)
        for r0 in range(0, n, rows_per):
            buf = _export(index, r0, min(rows_per, n - r0))
            crc = zlib.crc32(buf, crc)
            f.write(buf)
        f.seek(0)
        f.write(_HDR.pack(MAGIC, d, n, int(row_offset), index.metric_type, 0 if index.storage == "f32" else 1,
                          int(index.normalize), crc & 0xFFFFFFFF))


def load_shard(path: str, device: int | None = None, verify: bool = True):
    """-> (IndexFlat, row_offset).  Raises ValueError on a bad magic, size or checksum."""
    with open(path, "rb") as f:
        hdr = f.read(_HDR.size)
        if len(hdr) != _HDR.size:
            raise ValueError(f"{path}: truncated header")
        magic, d, n, row_offset, metric, storage, normalized, crc_want = _HDR.unpack(hdr)
        if magic != MAGIC:
            raise ValueError(f"{path}: not a b2r shard file (magic {magic!r})")
        if d <= 0 or n < 0 or metric not in (0, 1) or storage not in (0, 1):
            raise ValueError(f"{path}: corrupt header")
        esz = 4 if storage == 0 else 2
        index = IndexFlat(d, metric, storage="f32" if storage == 0 else "bf16", device=device,
                          normalize=bool(normalized))
        index.reserve(n)
        rows_per = max(1, _CHUNK_BYTES // (d * esz))
        crc = 0
        for r0 in range(0, n, rows_per):
            m = min(rows_per, n - r0)
            buf = np.frombuffer(f.read(m * d * esz), dtype=np.uint8)
            if buf.size != m * d * esz:
                raise ValueError(f"{path}: truncated payload")
            crc = zlib.crc32(buf, crc)
            check(_lib.lib().b2r_index_import_rows(index._h, m, buf.ctypes.data, 0))
        if verify and (crc & 0xFFFFFFFF) != crc_want:
            raise ValueError(f"{path}: payload checksum mismatch")
    return index, row_offset


# --------------------------------------------------------------------------------------------------
# faiss IndexFlat file layout (restated; see the module docstring)
# --------------------------------------------------------------------------------------------------
def _fourcc(s: str) -> int:
    return s.encode()[0] | (s.encode()[1] << 8) | (s.encode()[2] << 16) | (s.encode()[3] << 24)


def write_faiss_flat(index: IndexFlat, path: str) -> None:
    if index.storage != "f32":
        raise ValueError("a faiss flat index stores float32 rows: use an exact-mode index")
    n, d = index.ntotal, index.d
    with open(path, "wb") as f:
        f.write(struct.pack("<I", _fourcc("IxFI" if index.metric_type == METRIC_INNER_PRODUCT else "IxF2")))
        f.write(struct.pack("<iqqq?i", d, n, 1 << 20, 1 << 20, True, index.metric_type))
        f.write(struct.pack("<Q", n * d))
        rows_per = max(1, _CHUNK_BYTES // (d * 4))
        for r0 in range(0, n, rows_per):
            f.write(_export(index, r0, min(rows_per, n - r0)))


def read_faiss_flat(path: str, device: int | None = None) -> IndexFlat:
    with open(path, "rb") as f:
        (cc,) = struct.unpack("<I", f.read(4))
        names = {_fourcc("IxFI"): METRIC_INNER_PRODUCT, _fourcc("IxF2"): METRIC_L2, _fourcc("IxFl"): None}
        if cc not in names:
            raise ValueError(f"{path}: not a faiss IndexFlat file (fourcc {cc:#x})")
        d, n, _, _, _trained, metric = struct.unpack("<iqqq?i", f.read(struct.calcsize("<iqqq?i")))
        if metric not in (0, 1):
            raise ValueError(f"{path}: metric type {metric} is not supported (inner product and L2 only)")
        (count,) = struct.unpack("<Q", f.read(8))
        if count != n * d:
            raise ValueError(f"{path}: vector size {count} does not match ntotal*d = {n * d}")
        index = IndexFlat(d, metric, device=device)
        index.reserve(n)
        rows_per = max(1, _CHUNK_BYTES // (d * 4))
        for r0 in range(0, n, rows_per):
            m = min(rows_per, n - r0)
            buf = np.frombuffer(f.read(m * d * 4), dtype=np.float32)
            if buf.size != m * d:
                raise ValueError(f"{path}: truncated payload")
            index.add(buf.reshape(m, d))
    return index

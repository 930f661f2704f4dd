"""Shard save/load: CPU checks of the file layout, GPU round trips."""
import struct

import numpy as np
import pytest

import rag_tl_domainllm_optimizer_b200 as b2r
from rag_tl_domainllm_optimizer_b200 import shard_io


def test_header_layout_and_rejects_garbage(tmp_path):
    assert shard_io._HDR.size == 64
    p = tmp_path / "bad.b2r"
    p.write_bytes(b"NOTASHRD" + b"\0" * 56)
    with pytest.raises(ValueError, match="not a b2r shard"):
        shard_io.load_shard(str(p))
    p.write_bytes(b"B2R")
    with pytest.raises(ValueError, match="truncated header"):
        shard_io.load_shard(str(p))
    q = tmp_path / "bad.faiss"
    q.write_bytes(struct.pack("<I", 0x12345678) + b"\0" * 64)
    with pytest.raises(ValueError, match="not a faiss IndexFlat"):
        shard_io.read_faiss_flat(str(q))
    assert shard_io._fourcc("IxFI") == int.from_bytes(b"IxFI", "little")


@pytest.mark.gpu
@pytest.mark.parametrize("storage", ["f32", "bf16"])
@pytest.mark.parametrize("metric", [b2r.METRIC_INNER_PRODUCT, b2r.METRIC_L2])
def test_shard_round_trip_is_bit_identical(oracle, tmp_path, storage, metric):
    rng = np.random.default_rng(12)
    x = rng.standard_normal((5003, 100)).astype(np.float32)      # d not a multiple of the row pitch
    q = rng.standard_normal((9, 100)).astype(np.float32)
    a = b2r.IndexFlat(100, metric, storage=storage, normalize=True)
    a.add(x)
    path = str(tmp_path / "shard0.b2r")
    shard_io.save_shard(a, path, row_offset=70_000)
    b, off = shard_io.load_shard(path)
    assert off == 70_000 and b.ntotal == a.ntotal and b.storage == storage and b.metric_type == metric
    Da, Ia = a.search(q, 10)
    Db, Ib = b.search(q, 10)
    assert np.array_equal(Ia, Ib) and np.array_equal(Da.view(np.uint32), Db.view(np.uint32))
    assert np.array_equal(a.reconstruct(4321), b.reconstruct(4321))
    raw = bytearray(open(path, "rb").read())
    raw[100] ^= 0xFF                                                 # flip a payload byte
    open(path, "wb").write(bytes(raw))
    with pytest.raises(ValueError, match="checksum"):
        shard_io.load_shard(path)


@pytest.mark.gpu
def test_faiss_flat_file_round_trip(oracle, tmp_path):
    rng = np.random.default_rng(13)
    x = rng.standard_normal((2000, 64)).astype(np.float32)
    q = rng.standard_normal((5, 64)).astype(np.float32)
    a = b2r.IndexFlatL2(64)
    a.add(x)
    path = str(tmp_path / "flat.index")
    shard_io.write_faiss_flat(a, path)
    raw = open(path, "rb").read()
    assert raw[:4] == b"IxF2" and len(raw) == 4 + 33 + 8 + 2000 * 64 * 4
    assert struct.unpack("<i", raw[4:8])[0] == 64 and struct.unpack("<q", raw[8:16])[0] == 2000
    assert np.array_equal(np.frombuffer(raw[45:], dtype=np.float32).reshape(2000, 64), x)
    b = shard_io.read_faiss_flat(path)
    Da, Ia = a.search(q, 7)
    Db, Ib = b.search(q, 7)
    assert np.array_equal(Ia, Ib) and np.array_equal(Da.view(np.uint32), Db.view(np.uint32))
    D0, I0 = oracle.search_defined(x, q, 7, 1)
    assert np.array_equal(Ib, I0) and np.array_equal(Db.view(np.uint32), D0.view(np.uint32))

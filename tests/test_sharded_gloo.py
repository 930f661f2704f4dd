"""world_size-2 gloo test (CPU) of the multi-GPU host logic: partitioning, the single all-gather of packed
shard-local lists, and the merge.  The device kernels are replaced by the oracle here (the GPU versions are
covered by the -m gpu tests); what is exercised is ShardedIndexFlat's own code path."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _pack(D, I, metric):
    o = D.view(np.uint32).astype(np.uint64)
    o = np.where(o & 0x80000000, ~o & 0xFFFFFFFF, o | 0x80000000)
    if metric == 1:
        o = ~o & 0xFFFFFFFF
    key = (o << np.uint64(32)) | ((~I.astype(np.uint64)) & np.uint64(0xFFFFFFFF))
    return np.where(I >= 0, key, np.uint64(0))


def _unpack(keys, metric):
    o = (keys >> np.uint64(32)).astype(np.uint64)
    if metric == 1:
        o = ~o & np.uint64(0xFFFFFFFF)
    u = np.where(o & np.uint64(0x80000000), o & np.uint64(0x7FFFFFFF), ~o & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    ids = ((~keys) & np.uint64(0xFFFFFFFF)).astype(np.int64)
    return u.view(np.float32), ids


class _OracleLocal:
    """IndexFlat stand-in: shard-local search by the oracle, packed like b2r_index_search_packed."""

    def __init__(self, d, metric):
        self.d, self.metric, self.x = d, metric, np.zeros((0, d), dtype=np.float32)

    @property
    def ntotal(self):
        return self.x.shape[0]

    def add(self, x):
        self.x = np.concatenate([self.x, x])

    def search_packed(self, q, k, row_offset):
        from oracle import flat_oracle as fo
        D, I = fo.search_defined(self.x, q, k, self.metric)
        I = np.where(I >= 0, I + row_offset, -1)
        return torch.from_numpy(_pack(D, I, self.metric).view(np.int64))


def _merge(gathered, metric):
    g = gathered.numpy().view(np.uint64)
    G, nq, k = g.shape
    flat = np.transpose(g, (1, 0, 2)).reshape(nq, G * k)
    top = -np.sort(-flat.astype(np.uint64).view(np.int64) ^ np.int64(-2 ** 63), axis=1)  # unsigned order via bias
    top = (top ^ np.int64(-2 ** 63)).view(np.uint64)[:, :k]
    D, I = _unpack(top, metric)
    empty = top == 0
    return np.where(empty, -3.4028235e38 if metric == 0 else 3.4028235e38, D).astype(np.float32), np.where(empty, -1, I)


def _worker(rank, world, port, metric, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import rag_tl_domainllm_optimizer_b200.sharded as sh
    from oracle import flat_oracle as fo
    rng = np.random.default_rng(11)
    n, d, k = 1001, 32, 6
    x = rng.standard_normal((n, d)).astype(np.float32)
    x[900:905] = x[3:8]                                   # ties across the shard boundary
    q = np.concatenate([rng.standard_normal((5, d)).astype(np.float32), x[3:5]])
    lo, hi = sh.shard_bounds(n, world, rank)
    idx = sh.ShardedIndexFlat(d, metric, _local=_OracleLocal(d, metric), _merge=_merge)
    idx.add_local(x[lo:hi], lo, n)
    D, I = idx.search(q, k)
    D0, I0 = fo.search_defined(x, q, k, metric)
    ok = bool(np.array_equal(I, I0) and np.array_equal(D, D0))
    gathered = [None] * world
    dist.all_gather_object(gathered, ok)
    if rank == 0:
        out.put(all(gathered))
    dist.destroy_process_group()


@pytest.mark.parametrize("metric", [0, 1])
def test_sharded_search_world2_gloo(metric):
    from oracle import flat_oracle as fo
    fo.build()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + metric
    procs = [ctx.Process(target=_worker, args=(r, 2, port, metric, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert out.get(timeout=5) is True

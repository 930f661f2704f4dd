"""Row-sharded search on real GPUs (needs >= 2 devices; skipped otherwise): one process per GPU, the library's own
NCCL all-gather (`b2r_index_search_sharded`), result compared on every rank with the unsharded oracle — ids and score
bits — including ties that straddle the shard boundary and an unprovable query that takes the device-side second
pass in one shard only.  The same check through the `torch.distributed` route (`native=False`)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import rag_tl_domainllm_optimizer_b200 as b2r
    from oracle import flat_oracle as fo
    ok = True
    for storage, metric, d, k in (("f32", 0, 256, 10), ("bf16", 1, 128, 7)):
        rng = np.random.default_rng(17)                         # the same data on every rank
        n, nq = 60_001, 70
        x = fo.normalize(rng.standard_normal((n, d)).astype(np.float32))
        per = -(-n // world)
        x[per - 2:per + 3] = x[11]                              # one row five times, across the first shard boundary
        x[200:260] = x[12]                                      # 60 copies inside shard 0: its first pass cannot prove query 1
        q = fo.normalize(rng.standard_normal((nq, d)).astype(np.float32))
        q[0], q[1] = x[11], x[12]
        xo, qo = (fo.round_bf16(x), fo.round_bf16(q)) if storage == "bf16" else (x, q)
        D0, I0 = fo.search_ref(xo, qo, k, metric, extra=128)
        lo, hi = b2r.shard_bounds(n, world, rank)
        for native in (True, False):
            idx = b2r.ShardedIndexFlat(d, metric, storage=storage, device=rank, native=native)
            idx.add_local(x[lo:hi], lo, n)
            for xq in (q, torch.from_numpy(q).cuda()):          # host buffers and device tensors
                D, I = idx.search(xq, k)
                D = D.cpu().numpy() if hasattr(D, "cpu") else D
                I = I.cpu().numpy() if hasattr(I, "cpu") else I
                ok = ok and bool(np.array_equal(I, I0) and np.array_equal(D.view(np.uint32), D0.view(np.uint32)))
            if rank == 0:
                ok = ok and idx.local.last_stats()["fallback_queries"] >= 1 and idx.local.last_stats()["unverified"] == 0
            del idx
    flags = [None] * world
    dist.all_gather_object(flags, ok)
    if rank == 0:
        out.put(all(flags))
    dist.destroy_process_group()


def test_sharded_search_through_the_library_nccl():
    import torch
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    from oracle import flat_oracle as fo
    fo.build()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29600 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
        assert p.exitcode == 0
    assert out.get(timeout=5) is True

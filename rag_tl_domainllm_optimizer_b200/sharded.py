"""Row-sharded flat index over the GPUs of one box: one process per GPU, one all-gather per search.

Rank r owns the contiguous rows ``[r*ceil(N/G), min(N,(r+1)*ceil(N/G)))``; a search runs
the shard-local kernels with ``row_offset`` = first owned row, leaves the shard's top-k PACKED (one uint64
per hit), all-gathers the ``[nq,k]`` payloads with a single collective (NCCL over NVLink on GPUs) and merges
the ``G*k`` candidates per query with the same comparator — so the result equals the unsharded search.
"""
from __future__ import annotations

from typing import Tuple

import numpy as np

from ._lib import METRIC_INNER_PRODUCT
from .index import IndexFlat, merge_packed, merge_packed_strided


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    per = -(-n // world) if world > 0 else n
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


class ShardedIndexFlat:
    """Every rank constructs it, adds ITS rows (``add_local``) and calls ``search`` with the same queries."""

    def __init__(self, d: int, metric: int = METRIC_INNER_PRODUCT, *, storage: str = "f32", device: int | None = None,
                 normalize: bool = False, shadow: bool = False, group=None, _local=None, _merge=None):
        import torch.distributed as dist
        self._dist = dist
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        # _local / _merge let the CPU (gloo) tests stand in for the device kernels; product code never sets them
        self.local = _local if _local is not None else IndexFlat(d, metric, storage=storage, device=device, normalize=normalize, shadow=shadow)
        self._merge = _merge if _merge is not None else merge_packed
        self.d, self.metric_type = d, metric
        self.row_offset = 0
        self.ntotal_global = 0

    def add_local(self, x_local, row_offset: int, ntotal_global: int) -> None:
        """Append this rank's rows; ``row_offset`` is the global id of its first row."""
        if self.local.ntotal == 0:
            self.row_offset = int(row_offset)
        self.local.add(x_local)
        self.ntotal_global = int(ntotal_global)

    def search(self, x, k: int):
        import torch
        if self.world == 1 and self.row_offset == 0 and hasattr(self.local, "search"):
            return self.local.search(x, k)                                # one shard: nothing to gather or merge
        if not hasattr(self.local, "search_packed_async") or len(x) > 8192:
            return self._search_synchronous(x, k)
        # Everything is enqueued before the host looks at anything: shard-local search (its count of unproven
        # queries rides as the last word of the payload), the single all-gather, the merge.
        n, k = len(x), int(k)
        buf = self.local.search_packed_async(x, k, self.row_offset)      # [n*k + 1] on this GPU
        gathered = torch.empty((self.world, n * k + 1), dtype=buf.dtype, device=buf.device)
        self._dist.all_gather_into_tensor(gathered, buf.view(1, -1), group=self.group)   # the single collective
        D, I = merge_packed_strided(gathered, n, k, self.metric_type)
        if self.local.storage == "f32" and int(gathered[:, n * k].max().item()) != 0:
            # some shard could not prove its top-k (exact mode): every rank saw the same flags, so every rank
            # repeats the search with the 3xTF32 variant
            from ._lib import ENGINE_FMA, ENGINE_UMMA3, B2RError
            previous = getattr(self.local, "engine", 0)
            try:
                self.local.set_engine(ENGINE_UMMA3)
                try:
                    D, I = self._search_synchronous(x, k)
                except B2RError:                      # a shape tcgen05 does not take: the FMA engine does
                    self.local.set_engine(ENGINE_FMA)
                    D, I = self._search_synchronous(x, k)
            finally:
                self.local.set_engine(previous)
        return D, I

    def _search_synchronous(self, x, k: int):
        import torch
        packed = self.local.search_packed(x, k, self.row_offset)          # [nq,k] on this GPU
        if self.world == 1:
            return self._merge(packed.unsqueeze(0), self.metric_type)
        gathered = torch.empty((self.world,) + tuple(packed.shape), dtype=packed.dtype, device=packed.device)
        # the single collective; the flat [G*nq, k] view is the concatenation form every backend accepts
        self._dist.all_gather_into_tensor(gathered.view(-1, packed.shape[-1]), packed.contiguous(), group=self.group)
        return self._merge(gathered, self.metric_type)

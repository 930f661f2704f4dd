"""Row-sharded flat index over the GPUs of one box: one process per GPU, one all-gather per search.

Rank r owns the contiguous rows ``[r*ceil(N/G), min(N,(r+1)*ceil(N/G)))``; a search runs
the shard-local kernels with ``row_offset`` = first owned row, leaves the shard's top-k PACKED (one uint64
per hit), all-gathers the ``[nq,k]`` payloads with a single collective (NCCL over NVLink on GPUs) and merges
the ``G*k`` candidates per query with the same comparator — so the result equals the unsharded search.

On GPUs the whole step lives behind the C ABI (``b2r_index_search_sharded``: local search, ``ncclAllGather``
into a buffer the handle keeps, merge kernel, all enqueued on the caller's stream with no host round trip);
this class only sets the communicator up (rank 0's ``ncclUniqueId`` travels through ``torch.distributed``).
The ``torch.distributed`` route below it is the same three steps spelled out in Python: it serves backends
other than NCCL (the CPU ``gloo`` tests, with the kernels stood in for) and ``native=False``.
"""
from __future__ import annotations

from typing import Tuple

from ._lib import METRIC_INNER_PRODUCT
from .index import IndexFlat, merge_packed, nccl_unique_id


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    per = -(-n // world) if world > 0 else n
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


class ShardedIndexFlat:
    """Every rank constructs it, adds ITS rows (``add_local``) and calls ``search`` with the same queries."""

    def __init__(self, d: int, metric: int = METRIC_INNER_PRODUCT, *, storage: str = "f32", device: int | None = None,
                 normalize: bool = False, shadow: bool = False, group=None, native: bool = True, _local=None, _merge=None):
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
        self._want_native = bool(native) and _local is None
        self._native_ready = False
        self._gather_buf = None

    def add_local(self, x_local, row_offset: int, ntotal_global: int) -> None:
        """Append this rank's rows; ``row_offset`` is the global id of its first row."""
        if self.local.ntotal == 0:
            self.row_offset = int(row_offset)
            self._native_ready = False
        self.local.add(x_local)
        self.ntotal_global = int(ntotal_global)

    def _setup_native(self) -> bool:
        """Collective, once: every rank joins the library's own NCCL communicator."""
        if self._native_ready:
            return True
        if not (self._want_native and self.world > 1 and self._dist.get_backend(self.group) == "nccl"):
            return False
        box = [nccl_unique_id() if self.rank == 0 else None]
        src = self._dist.get_global_rank(self.group, 0) if self.group is not None else 0
        self._dist.broadcast_object_list(box, src=src, group=self.group)
        self.local.comm_init(box[0], self.rank, self.world, self.row_offset)
        self._native_ready = True
        return True

    def search(self, x, k: int):
        if self.world == 1 and self.row_offset == 0 and hasattr(self.local, "search"):
            return self.local.search(x, k)                                # one shard: nothing to gather or merge
        if self._setup_native():
            return self.local.search_sharded(x, k)                        # search + ncclAllGather + merge, in the library
        return self._search_torch(x, k)

    def _search_torch(self, x, k: int):
        import torch
        packed = self.local.search_packed(x, k, self.row_offset)          # [nq,k] on this rank's device
        if self.world == 1:
            return self._merge(packed.unsqueeze(0), self.metric_type)
        shape = (self.world,) + tuple(packed.shape)
        if self._gather_buf is None or tuple(self._gather_buf.shape) != shape or self._gather_buf.device != packed.device:
            self._gather_buf = torch.empty(shape, dtype=packed.dtype, device=packed.device)   # kept across searches
        # the single collective; the flat [G*nq, k] view is the concatenation form every backend accepts
        self._dist.all_gather_into_tensor(self._gather_buf.view(-1, packed.shape[-1]), packed.contiguous(), group=self.group)
        return self._merge(self._gather_buf, self.metric_type)

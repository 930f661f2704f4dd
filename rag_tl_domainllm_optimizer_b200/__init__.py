"""B200-native exact dense retrieval: the FAISS-flat-shaped retrieval step of the reference's RAG loop.

Public surface (mirrors ``faiss``): ``IndexFlatIP``, ``IndexFlatL2``, ``IndexFlat``, ``normalize_L2``,
``METRIC_INNER_PRODUCT``, ``METRIC_L2``; plus ``ShardedIndexFlat`` (row-sharded over the GPUs of one box)
and ``FlatRetriever`` (id -> passage text, producing the ``retrieved_docs`` lists the reference consumes).
"""
from ._lib import (ENGINE_AUTO, ENGINE_FMA, ENGINE_SHADOW, ENGINE_UMMA, ENGINE_UMMA3, METRIC_INNER_PRODUCT, METRIC_L2, B2RError)
from .index import (IndexFlat, IndexFlatIP, IndexFlatL2, ShapeError, cosine_rows, merge_packed, merge_packed_strided,
                    normalize_L2)
from .retriever import FlatRetriever, collate_retrieval_batch, raft_select
from .sharded import ShardedIndexFlat, shard_bounds

__all__ = [
    "IndexFlat", "IndexFlatIP", "IndexFlatL2", "normalize_L2", "METRIC_INNER_PRODUCT", "METRIC_L2", "ShapeError",
    "B2RError", "ShardedIndexFlat", "shard_bounds", "FlatRetriever", "collate_retrieval_batch", "raft_select", "merge_packed", "merge_packed_strided",
    "cosine_rows", "ENGINE_AUTO", "ENGINE_FMA", "ENGINE_UMMA", "ENGINE_UMMA3", "ENGINE_SHADOW",
]

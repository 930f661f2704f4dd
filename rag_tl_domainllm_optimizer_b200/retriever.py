"""Retriever adapter: embeddings in, passages out — the producer of ``retrieved_docs``.

The reference reads retrieved passages from a dataset column
(reinforcement_learning_optimization_after_rag.py:286-288, :405, literal at :507) and feeds them to
``RAGEnvironment.generate_response(query, retrieved_docs)`` (:31-34) and
``RewardModel.calculate_factual_accuracy(response, retrieved_docs)`` (:63-71), both of which iterate
``for doc in retrieved_docs`` — i.e. they expect a ``list[str]`` per query, in rank order.  This class
is the LangChain-``FAISS``-store-shaped piece that produces exactly that from an index search:
``index_to_docstore_id`` + docstore, ``-1`` ids skipped.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

import numpy as np

from .index import IndexFlat


class FlatRetriever:
    def __init__(self, index: IndexFlat, passages: Sequence[str] | None = None):
        self.index = index
        self.passages: List[str] = list(passages) if passages is not None else []

    def add(self, embeddings, passages: Iterable[str]) -> None:
        passages = list(passages)
        if len(passages) != len(embeddings):
            raise ValueError(f"{len(embeddings)} embeddings for {len(passages)} passages")
        self.index.add(embeddings)
        self.passages.extend(passages)

    def search_with_scores(self, query_embeddings, k: int) -> List[List[Tuple[str, float]]]:
        D, I = self.index.search(np.ascontiguousarray(query_embeddings, dtype=np.float32), k)
        out = []
        for drow, irow in zip(np.asarray(D), np.asarray(I)):
            out.append([(self.passages[int(i)], float(s)) for s, i in zip(drow, irow) if i >= 0])
        return out

    def retrieve(self, query_embeddings, k: int) -> List[List[str]]:
        """-> one ``list[str]`` of up to k passages per query, best first (the ``retrieved_docs`` contract)."""
        return [[p for p, _ in row] for row in self.search_with_scores(query_embeddings, k)]


def collate_retrieval_batch(items: Sequence[dict]) -> dict:
    """``collate_fn`` that keeps ``retrieved_docs`` as one list[str] per sample.

    torch's default_collate transposes a list column into k tuples of B strings, which makes
    ``zip(queries, retrieved_docs_batch)`` at reference :292 / :299 pair query i with the i-th ranked
    passage of every query.  Use this with ``DataLoader(..., collate_fn=...)``.
    """
    keys = items[0].keys() if items else []
    return {key: [it[key] for it in items] for key in keys}

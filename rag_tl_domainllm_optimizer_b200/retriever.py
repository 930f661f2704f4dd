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

    def retrieve_raft(self, query_embeddings, gold_ids, k: int, k_prime: int | None = None, p_gold: float = 1.0,
                      seed: int = 0) -> List[List[str]]:
        """Gold + distractor passages per query (RAFT rollout sets): search top-k', then ``raft_select``."""
        k_prime = k_prime or 2 * k
        _, I = self.index.search(np.ascontiguousarray(query_embeddings, dtype=np.float32), k_prime)
        ids, _ = raft_select(np.asarray(I), gold_ids, k, p_gold, seed)
        return [[self.passages[int(j)] for j in row if j >= 0] for row in ids]

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


def raft_select(I, gold_ids, k: int, p_gold: float = 1.0, seed: int = 0):
    """RAFT-style document sets from a top-k' search result (reference README.md:2, config 4).

    For each query: the gold document (``gold_ids[i]``; kept with probability ``p_gold``, as RAFT trains a share
    of samples without it) plus distractors = the best-ranked retrieved ids that are not the gold one, ``k`` ids in
    all, then shuffled so that the gold position carries no signal.  ``I`` is the ``[n, k']`` id matrix of a search
    with k' >= k (``-1`` = no hit).  Returns (ids [n, k] int64 with -1 padding, gold_pos [n] int64, -1 if absent).
    Pure index bookkeeping on the search output — the search itself is the GPU path.
    """
    I = np.asarray(I)
    n = I.shape[0]
    rng = np.random.default_rng(seed)
    out = np.full((n, k), -1, dtype=np.int64)
    gold_pos = np.full(n, -1, dtype=np.int64)
    for i in range(n):
        gold = int(gold_ids[i])
        keep_gold = gold >= 0 and rng.random() < p_gold
        distractors = [int(j) for j in I[i] if j >= 0 and int(j) != gold]
        docs = ([gold] if keep_gold else []) + distractors[: k - (1 if keep_gold else 0)]
        order = rng.permutation(len(docs))
        docs = [docs[j] for j in order]
        out[i, : len(docs)] = docs
        if keep_gold:
            gold_pos[i] = docs.index(gold)
    return out, gold_pos

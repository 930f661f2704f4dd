"""CPU tests of the host-side logic that needs no device: sharding bounds, retriever adapter, collation."""
import numpy as np
import pytest

import rag_tl_domainllm_optimizer_b200 as b2r


def test_shard_bounds_cover_and_are_contiguous():
    for n in (0, 1, 7, 8, 9, 1000, 10_000_000):
        for g in (1, 2, 4, 8):
            spans = [b2r.shard_bounds(n, g, r) for r in range(g)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
                assert a1 == b0 and a0 <= a1
            per = -(-n // g) if n else 0
            assert all(hi - lo <= per for lo, hi in spans)


class _FakeIndex:
    """Stands in for IndexFlat so the adapter logic can run without a GPU."""

    def __init__(self, D, I):
        self.D, self.I, self.added = D, I, 0

    def add(self, x):
        self.added += len(x)

    def search(self, x, k):
        return self.D[:, :k], self.I[:, :k]


def test_retriever_emits_list_of_str_per_query_and_skips_missing():
    D = np.array([[0.9, 0.5, -3.4e38], [0.7, 0.1, 0.0]], dtype=np.float32)
    I = np.array([[2, 0, -1], [1, 2, 0]], dtype=np.int64)
    r = b2r.FlatRetriever(_FakeIndex(D, I))
    r.add(np.zeros((3, 4), dtype=np.float32), ["doc a", "doc b", "doc c"])
    docs = r.retrieve(np.zeros((2, 4), dtype=np.float32), 3)
    assert docs == [["doc c", "doc a"], ["doc b", "doc c", "doc a"]]       # -1 skipped, rank order kept
    # exactly what reference :33 / :67 iterate: `for doc in retrieved_docs` yields passages, not characters
    assert all(isinstance(d, str) and len(d) > 1 for row in docs for d in row)
    with pytest.raises(ValueError):
        r.add(np.zeros((2, 4), dtype=np.float32), ["only one"])


def test_collate_keeps_per_sample_lists():
    items = [{"query": "q0", "retrieved_docs": ["a", "b"]}, {"query": "q1", "retrieved_docs": ["c", "d"]},
             {"query": "q2", "retrieved_docs": ["e", "f"]}]
    batch = b2r.collate_retrieval_batch(items)
    pairs = list(zip(batch["query"], batch["retrieved_docs"]))              # the zip at reference :292
    assert pairs == [("q0", ["a", "b"]), ("q1", ["c", "d"]), ("q2", ["e", "f"])]


def test_raft_select_gold_plus_distractors():
    I = np.array([[5, 9, 2, 7, 1, 3], [4, 8, 6, 0, -1, -1], [1, 2, 3, 4, 5, 6]], dtype=np.int64)
    gold = [2, 33, 6]                      # in the hits / not retrieved at all / last of the hits
    ids, pos = b2r.raft_select(I, gold, k=4, p_gold=1.0, seed=3)
    for row, g, p in zip(ids, gold, pos):
        assert row[p] == g and list(row).count(g) == 1            # gold present exactly once, position reported
    assert sorted(ids[0]) == [2, 5, 7, 9]                          # gold + the 3 best-ranked non-gold hits
    assert sorted(ids[1]) == [4, 6, 8, 33]
    assert sorted(ids[2]) == [1, 2, 3, 6]
    ids0, pos0 = b2r.raft_select(I, gold, k=4, p_gold=0.0, seed=3)  # RAFT's "no gold" share: distractors only
    assert np.all(pos0 == -1) and sorted(ids0[0]) == [5, 7, 9, 1] or sorted(ids0[0]) == [1, 5, 7, 9]
    assert 2 not in ids0[0]
    a, _ = b2r.raft_select(I, gold, 4, 1.0, seed=11)
    b, _ = b2r.raft_select(I, gold, 4, 1.0, seed=11)
    assert np.array_equal(a, b)                                    # deterministic in the seed
    r = b2r.FlatRetriever(_FakeIndex(np.zeros((3, 6), np.float32), I), [f"p{i}" for i in range(40)])
    docs = r.retrieve_raft(np.zeros((3, 4), np.float32), gold, k=4, k_prime=6, seed=3)
    assert [len(d) for d in docs] == [4, 4, 4] and "p2" in docs[0] and "p33" in docs[1]


def test_oracle_rescore_rows_orders_by_global_id(oracle):
    """bench.py's full-size parity check hands the oracle candidate rows with their own global ids."""
    rng = np.random.default_rng(5)
    x = rng.standard_normal((400, 48)).astype(np.float32)
    x[300] = x[9]                                      # a tie: the lower GLOBAL id must come first
    q = x[9].copy()
    D0, I0 = oracle.search_defined(x, q[None], 6, 0)
    sel = np.unique(np.concatenate([I0[0], rng.integers(0, 400, 40)]))[::-1].copy()     # any order
    D, I = oracle.rescore_rows(x[sel], sel, q, 6, 0)
    assert np.array_equal(I, I0[0]) and np.array_equal(D.view(np.uint32), D0[0].view(np.uint32))
    assert list(I[:2]) == [9, 300]


def test_bench_rows_do_not_depend_on_the_sharding():
    """bench.py generates the corpus per 65536-row block: any shard of it sees the same rows (all distributions)."""
    import torch
    import bench
    dev = torch.device("cpu")
    for dist in ("gauss", "mixture", "dup"):
        whole = bench.gen_rows(torch, 0, 65000, 1500, 32, dev, dist)
        parts = torch.cat([bench.gen_rows(torch, 0, 65000, 700, 32, dev, dist), bench.gen_rows(torch, 0, 65700, 800, 32, dev, dist)])
        assert torch.equal(whole, parts)
    dup = bench.gen_rows(torch, 0, 0, 200, 32, dev, "dup")
    assert torch.equal(dup[99], dup[49]) and torch.equal(dup[199], dup[149])       # 1 % exact duplicates
    q = bench.gen_queries(torch, 8, 32, dev, "mixture", 5000)
    assert q.shape == (8, 32)


def test_bench_reads_traffic_from_the_committed_ncu_capture():
    import bench
    traffic, src = bench.ncu_traffic("exact", 1024, bench.N_ROWS, 1)
    assert src and src.startswith("profiles/") and 3.0e9 < traffic < 3.6e9          # ~1.05x the 3.07 GB corpus
    assert bench.ncu_traffic("exact", 1024, 12345, 1) == (None, None)

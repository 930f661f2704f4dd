"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle.

Bar (BASELINE.json north_star): exact mode — ids identical to the oracle (ties: lower id first) and scores
within 1e-5 relative; here the engine's arithmetic is fully defined, so both ids AND fp32
scores are required to be BIT-IDENTICAL to oracle/flat_oracle.c.  bf16 mode — bit-identical to the oracle
run on bf16-rounded inputs; recall@k against the fp32 oracle is reported by bench.py.
"""
import ctypes
import os

import numpy as np
import pytest

import rag_tl_domainllm_optimizer_b200 as b2r
from rag_tl_domainllm_optimizer_b200 import _lib

pytestmark = pytest.mark.gpu
FLT_MAX = float(np.finfo(np.float32).max)
ENGINES = [b2r.ENGINE_FMA, b2r.ENGINE_AUTO]


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _check_exact(D, I, D0, I0):
    assert np.array_equal(np.asarray(I), I0), f"id mismatch in {np.argwhere(np.asarray(I) != I0)[:5]}"
    assert np.array_equal(_bits(D), _bits(D0))


def _mk(oracle, n, d, nq, seed, normalize=True):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, d)).astype(np.float32)
    q = rng.standard_normal((nq, d)).astype(np.float32)
    if normalize:
        x, q = oracle.normalize(x), oracle.normalize(q)
    return x, q


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("metric", [b2r.METRIC_INNER_PRODUCT, b2r.METRIC_L2])
def test_config1_minilm_10k(oracle, engine, metric):
    """BASELINE.json configs[0]: 10k x 384, 64 queries, k=5, exact."""
    x, q = _mk(oracle, 10_000, 384, 64, 100)
    idx = b2r.IndexFlat(384, metric)
    idx.set_engine(engine)
    idx.add(x)
    assert idx.ntotal == 10_000
    D, I = idx.search(q, 5)
    D0, I0 = oracle.search_defined(x, q, 5, metric)
    _check_exact(D, I, D0, I0)
    st = idx.last_stats()
    assert st["unverified"] == 0 and st["kprime"] >= 5


@pytest.mark.parametrize("engine", ENGINES)
def test_golden_search_vectors(oracle, golden_dir, engine):
    g = np.load(os.path.join(golden_dir, "search_small.npz"))
    for ci in range(int(g["ncases"])):
        k, metric = (int(v) for v in g[f"c{ci}_meta"])
        x, q = g[f"c{ci}_x"], g[f"c{ci}_q"]
        idx = b2r.IndexFlat(x.shape[1], metric)
        idx.set_engine(engine)
        idx.add(x)
        D, I = idx.search(q, k)
        _check_exact(D, I, g[f"c{ci}_D"], g[f"c{ci}_I"])


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("n,d,nq,k", [(1, 8, 1, 1), (33, 100, 3, 7), (1000, 4, 9, 10), (4097, 260, 17, 33),
                                      (2500, 1024, 5, 100), (700, 4096, 2, 4), (5000, 64, 70, 1)])
def test_ragged_shapes(oracle, engine, n, d, nq, k):
    x, q = _mk(oracle, n, d, nq, n + d, normalize=False)
    for metric in (b2r.METRIC_INNER_PRODUCT, b2r.METRIC_L2):
        idx = b2r.IndexFlat(d, metric)
        idx.set_engine(engine)
        idx.add(x)
        D, I = idx.search(q, k)
        D0, I0 = oracle.search_defined(x, q, k, metric)
        _check_exact(D, I, D0, I0)


@pytest.mark.parametrize("engine", ENGINES)
def test_ties_padding_empty(oracle, engine):
    rng = np.random.default_rng(4)
    base = rng.standard_normal((500, 96)).astype(np.float32)
    x = np.concatenate([base, base, base[:100]])
    q = base[:9].copy()
    for metric in (b2r.METRIC_INNER_PRODUCT, b2r.METRIC_L2):
        idx = b2r.IndexFlat(96, metric)
        idx.set_engine(engine)
        D, I = idx.search(q, 3)                                        # empty index
        assert np.all(I == -1) and np.all(D == (-FLT_MAX if metric == 0 else FLT_MAX))
        idx.add(x[:600])
        idx.add(x[600:])                                               # second add: ids continue
        assert idx.ntotal == 1100
        D, I = idx.search(q, 8)
        D0, I0 = oracle.search_defined(x, q, 8, metric)
        _check_exact(D, I, D0, I0)
        assert np.all(I[:, 0] + 500 == I[:, 1])                        # exact duplicates: lower id first
        small = b2r.IndexFlat(96, metric)
        small.set_engine(engine)
        small.add(x[:4])
        D, I = small.search(q, 7)                                      # k > ntotal
        D0, I0 = oracle.search_defined(x[:4], q, 7, metric)
        _check_exact(D, I, D0, I0)
        assert np.all(I[:, 4:] == -1)
        idx.reset()
        assert idx.ntotal == 0
        idx.add(x[:10])
        D, I = idx.search(q, 2)
        _check_exact(D, I, *oracle.search_defined(x[:10], q, 2, metric))


@pytest.mark.parametrize("engine", ENGINES)
def test_bf16_mode_equals_oracle_on_rounded_inputs(oracle, engine):
    x, q = _mk(oracle, 20_000, 768, 40, 77)
    xr, qr = oracle.round_bf16(x), oracle.round_bf16(q)
    for metric in (b2r.METRIC_INNER_PRODUCT, b2r.METRIC_L2):
        idx = b2r.IndexFlat(768, metric, storage="bf16")
        idx.set_engine(engine)
        idx.add(x)
        D, I = idx.search(q, 10)
        D0, I0 = oracle.search_defined(xr, qr, 10, metric)
        _check_exact(D, I, D0, I0)
        assert np.array_equal(idx.reconstruct(123), xr[123])


def test_normalize_on_ingest_and_queries(oracle):
    rng = np.random.default_rng(9)
    x = rng.standard_normal((3000, 384)).astype(np.float32) * rng.uniform(0.01, 100, (3000, 1)).astype(np.float32)
    x[17] = 0.0                                                        # zero-norm row stays zero
    q = rng.standard_normal((12, 384)).astype(np.float32) * 7
    idx = b2r.IndexFlatIP(384, normalize=True)
    idx.add(x)
    xn, qn = oracle.normalize(x), oracle.normalize(q)
    assert np.array_equal(_bits(idx.reconstruct(5)), _bits(xn[5]))
    assert np.all(idx.reconstruct(17) == 0)
    D, I = idx.search(q, 10)
    _check_exact(D, I, *oracle.search_defined(xn, qn, 10, 0))
    y = x.copy()
    b2r.normalize_L2(y)                                                # faiss.normalize_L2 replacement, in place
    assert np.array_equal(_bits(y), _bits(xn))
    y2 = np.ascontiguousarray(x[:, :101])                              # unaligned row pitch
    b2r.normalize_L2(y2)
    assert np.array_equal(_bits(y2), _bits(oracle.normalize(np.ascontiguousarray(x[:, :101]))))


def test_reward_cosine_golden_through_the_gpu(oracle, golden_dir):
    """The reference's own RewardModel outputs (sklearn cosine_similarity at :70/:78/:104) vs the GPU path."""
    g = np.load(os.path.join(golden_dir, "reward_cosine.npz"))
    for ci in range(int(g["ncases"])):
        resp, docs, query = g[f"c{ci}_resp"], g[f"c{ci}_docs"], g[f"c{ci}_query"]
        cos, mx = b2r.cosine_rows(resp, docs)
        cos0, mx0 = oracle.cosine_rows(resp, docs)
        assert np.array_equal(_bits(mx), _bits(mx0)) and np.array_equal(_bits(cos), _bits(cos0))
        np.testing.assert_allclose(mx, g[f"c{ci}_fact"], rtol=0, atol=2e-6)
        _, rel = b2r.cosine_rows(resp, query[:, None, :])
        np.testing.assert_allclose(rel, g[f"c{ci}_rel"], rtol=0, atol=2e-6)
        _, gts = b2r.cosine_rows(resp, g[f"c{ci}_gt"][:, None, :])     # ground-truth similarity, reference :99-104
        np.testing.assert_allclose(gts, g[f"c{ci}_gtsim"], rtol=0, atol=2e-6)
        if docs.shape[1]:
            for i in range(resp.shape[0]):                             # same number through IndexFlatIP, k=1
                idx = b2r.IndexFlatIP(resp.shape[1], normalize=True)
                idx.add(docs[i])
                D, _ = idx.search(resp[i:i + 1], 1)
                assert abs(float(D[0, 0]) - float(g[f"c{ci}_fact"][i])) <= 2e-6


def test_torch_cuda_tensors_and_streams(oracle):
    import torch
    x, q = _mk(oracle, 5000, 256, 21, 5)
    idx = b2r.IndexFlatIP(256)
    idx.add(torch.from_numpy(x).cuda())
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        D, I = idx.search(torch.from_numpy(q).cuda(), 10)
    s.synchronize()
    assert D.is_cuda and I.dtype == torch.int64
    _check_exact(D.cpu().numpy(), I.cpu().numpy(), *oracle.search_defined(x, q, 10, 0))
    Dc, Ic = idx.search(torch.from_numpy(q), 10)                       # CPU tensor in -> CPU tensors out
    assert not Dc.is_cuda
    _check_exact(Dc.numpy(), Ic.numpy(), *oracle.search_defined(x, q, 10, 0))


def test_shape_errors():
    idx = b2r.IndexFlatIP(16)
    with pytest.raises(AssertionError):
        idx.add(np.zeros((3, 15), dtype=np.float32))
    with pytest.raises(b2r.ShapeError):
        idx.search(np.zeros((3, 16), dtype=np.float64), 1)
    with pytest.raises(b2r.ShapeError):
        idx.search(np.zeros((3, 16), dtype=np.float32), 0)
    with pytest.raises(b2r.B2RError):
        idx.search(np.zeros((3, 16), dtype=np.float32), 5000)


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("storage", ["f32", "bf16"])
def test_mid_size_vs_ref(oracle, engine, storage):
    """200k x 768 (a slice of config 2's shape), batch 48, k=10, against search_ref."""
    x = oracle.normalize(oracle.synth_rows(0, 0, 200_000, 768))
    q = oracle.normalize(oracle.synth_rows(1, 0, 48, 768))
    idx = b2r.IndexFlatIP(768, storage=storage)
    idx.set_engine(engine)
    idx.add(x)
    D, I = idx.search(q, 10)
    if storage == "bf16":
        x, q = oracle.round_bf16(x), oracle.round_bf16(q)
    _check_exact(D, I, *oracle.search_ref(x, q, 10, 0))
    assert idx.last_stats()["unverified"] == 0


def test_sharded_equals_unsharded_on_one_gpu(oracle):
    """G shard handles on one device + the packed merge == one unsharded search (ties across shards)."""
    import torch
    x, q = _mk(oracle, 30_001, 128, 19, 8)
    x[20_000:20_050] = x[100:150]
    q[:4] = x[100:104]
    for metric in (b2r.METRIC_INNER_PRODUCT, b2r.METRIC_L2):
        D0, I0 = oracle.search_ref(x, q, 10, metric)
        for G in (2, 4, 8):
            packs = []
            for r in range(G):
                lo, hi = b2r.shard_bounds(len(x), G, r)
                sh = b2r.IndexFlat(128, metric)
                sh.add(x[lo:hi])
                packs.append(sh.search_packed(q, 10, lo))
            D, I = b2r.merge_packed(torch.stack(packs), metric)
            _check_exact(D.cpu().numpy(), I.cpu().numpy(), D0, I0)


def test_c_abi_direct_host_pointers(oracle):
    """The raw C ABI as a C caller would use it (host buffers in, host buffers out)."""
    L = _lib.lib()
    x, q = _mk(oracle, 2000, 64, 6, 3)
    h = ctypes.c_void_p()
    assert L.b2r_index_create(ctypes.byref(h), 64, 0, 0, 0) == 0
    assert L.b2r_index_add(h, 2000, x.ctypes.data, 0, None) == 0
    n = ctypes.c_int64()
    assert L.b2r_index_ntotal(h, ctypes.byref(n)) == 0 and n.value == 2000
    D = np.empty((6, 4), dtype=np.float32)
    I = np.empty((6, 4), dtype=np.int64)
    assert L.b2r_index_search(h, 6, q.ctypes.data, 4, D.ctypes.data, I.ctypes.data, None) == 0
    _check_exact(D, I, *oracle.search_defined(x, q, 4, 0))
    assert L.b2r_index_search(h, 6, q.ctypes.data, 0, D.ctypes.data, I.ctypes.data, None) == -1
    assert b"k must be positive" in L.b2r_last_error()
    L.b2r_index_free(h)


def test_retriever_over_a_real_index_emits_the_oracles_passages(oracle):
    """FlatRetriever over a real IndexFlatIP produces, per query, the list[str] that the reference
    consumes at :31-34 / :63-71 — the passages of the oracle's top-k ids, in rank order, with -1 padding skipped."""
    rng = np.random.default_rng(21)
    n, d, nq, k = 5000, 384, 33, 4
    x = oracle.normalize(rng.standard_normal((n, d)).astype(np.float32))
    x[4000:4003] = x[10]                                   # duplicates: the lower id comes first in the passages too
    q = oracle.normalize(rng.standard_normal((nq, d)).astype(np.float32))
    q[0] = x[10]
    passages = [f"passage #{i}: lorem {i * 7919 % 1000}" for i in range(n)]
    r = b2r.FlatRetriever(b2r.IndexFlatIP(d))
    r.add(x[:3000], passages[:3000])
    r.add(x[3000:], passages[3000:])
    D0, I0 = oracle.search_defined(x, q, k, 0)
    docs = r.retrieve(q, k)
    assert docs == [[passages[i] for i in row] for row in I0]
    assert docs[0][:2] == [passages[10], passages[4000]]
    scored = r.search_with_scores(q, k)
    assert [[p for p, _ in row] for row in scored] == docs
    assert np.array_equal(np.array([[s for _, s in row] for row in scored], dtype=np.float32).view(np.uint32), D0.view(np.uint32))
    few = b2r.FlatRetriever(b2r.IndexFlatIP(d))            # k > ntotal: the -1 slots yield no passage
    few.add(x[:2], passages[:2])
    assert [len(row) for row in few.retrieve(q[:3], 4)] == [2, 2, 2]
    # the collated batch pairs each query with ITS passages (the zip at reference :292 / :299)
    batch = b2r.collate_retrieval_batch([{"query": f"q{i}", "retrieved_docs": docs[i]} for i in range(nq)])
    assert list(zip(batch["query"], batch["retrieved_docs"]))[5] == ("q5", docs[5])


def test_raft_sets_over_a_real_index(oracle):
    """SURVEY §8 f4 (configs[3] shape, reduced): gold + distractor sets built from a real top-k' search equal the
    sets built from the oracle's ids."""
    rng = np.random.default_rng(22)
    n, d, nq, k = 3000, 4096, 16, 4
    x = oracle.normalize(rng.standard_normal((n, d)).astype(np.float32))
    gold = rng.integers(0, n, nq)
    q = oracle.normalize(x[gold] + 0.5 * oracle.normalize(rng.standard_normal((nq, d)).astype(np.float32)))
    gold[3] = -1                                           # a sample without a gold document
    passages = [f"doc {i}" for i in range(n)]
    r = b2r.FlatRetriever(b2r.IndexFlatIP(d, storage="bf16"), passages)
    r.index.add(x)
    sets = r.retrieve_raft(q, gold, k, k_prime=8, seed=5)
    _, I0 = oracle.search_ref(oracle.round_bf16(x), oracle.round_bf16(q), 8, 0)
    ids0, pos0 = b2r.raft_select(I0, gold, k, 1.0, 5)
    assert sets == [[passages[j] for j in row if j >= 0] for row in ids0]
    for i in range(nq):
        assert len(sets[i]) == k and (gold[i] < 0 or sets[i][pos0[i]] == passages[gold[i]])

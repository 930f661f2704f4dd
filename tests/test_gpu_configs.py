"""GPU parity at the shapes of BASELINE.json's other configs (reduced row counts so the oracle finishes in
seconds), the multi-launch path (more than 2048 queries), the proof/fallback path, and size-independent
properties at the full 1M x 768 size of configs[1]."""
import numpy as np
import pytest

import rag_tl_domainllm_optimizer_b200 as b2r

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _check(D, I, D0, I0):
    assert np.array_equal(np.asarray(I), I0)
    assert np.array_equal(_bits(D), _bits(D0))


def _data(oracle, n, d, nq, seed0):
    x = oracle.normalize(oracle.synth_rows(seed0, 0, n, d))
    q = oracle.normalize(oracle.synth_rows(seed0 + 1, 0, nq, d))
    return x, q


@pytest.mark.parametrize("storage", ["f32", "bf16"])
def test_config3_shape_bge_large_k100(oracle, storage):
    """configs[2] shape: d=1024, batch 256, k=100 (one shard's worth of rows, reduced)."""
    x, q = _data(oracle, 60_000, 1024, 256, 30)
    idx = b2r.IndexFlatIP(1024, storage=storage)
    idx.add(x)
    D, I = idx.search(q, 100)
    if storage == "bf16":
        x, q = oracle.round_bf16(x), oracle.round_bf16(q)
    _check(D, I, *oracle.search_ref(x, q, 100, 0, extra=100))
    st = idx.last_stats()
    assert st["engine"] == b2r.ENGINE_UMMA and st["unverified"] == 0


@pytest.mark.parametrize("storage", ["f32", "bf16"])
def test_config4_shape_llama_hidden_k4(oracle, storage):
    """configs[3] shape: d=4096 (LLaMA-2-7B hidden size), 256 queries per step, k=4 gold+distractor docs."""
    x, q = _data(oracle, 20_000, 4096, 256, 40)
    idx = b2r.IndexFlatIP(4096, storage=storage)
    idx.add(x)
    D, I = idx.search(q, 4)
    if storage == "bf16":
        x, q = oracle.round_bf16(x), oracle.round_bf16(q)
    _check(D, I, *oracle.search_ref(x, q, 4, 0))
    assert idx.last_stats()["unverified"] == 0


def test_config5_shape_batch4096_bf16(oracle):
    """configs[4] shape: d=768 bf16, batch 4096 (two query chunks of 2048 per scoring launch set), k=10."""
    x, q = _data(oracle, 100_000, 768, 4096, 50)
    idx = b2r.IndexFlatIP(768, storage="bf16")
    idx.add(x)
    D, I = idx.search(q, 10)
    _check(D, I, *oracle.search_ref(oracle.round_bf16(x), oracle.round_bf16(q), 10, 0))
    st = idx.last_stats()
    assert st["unverified"] == 0 and st["nq"] == 4096


def test_unproven_query_is_researched_alone(oracle):
    """60 exact copies of one row: the gap between the k-th and the K'-th candidate is zero, so the first pass
    cannot prove query 0.  Only that query goes through the range pass (every row above a fixed bound is
    re-scored: no gap needed) and the call still returns the lowest ids; the other 1023 queries of the batch
    are not searched again."""
    x, q = _data(oracle, 30_000, 256, 1024, 60)
    x[1000:1060] = x[7]
    q[0] = x[7]
    idx = b2r.IndexFlatIP(256)
    idx.add(x)
    D, I = idx.search(q, 10)
    _check(D, I, *oracle.search_ref(x, q, 10, 0, extra=128))
    assert I[0, 0] == 7 and list(I[0, 1:]) == list(range(1000, 1009))
    st = idx.last_stats()
    assert st["fallback_queries"] == 1 and st["exhaustive_queries"] == 0 and st["unverified"] == 0
    assert st["fallback"] == 1 and st["engine"] == b2r.ENGINE_UMMA
    idx.set_verify(False)                       # without the hand-over the query is only counted
    idx.search(q, 10)
    st = idx.last_stats()
    assert st["unverified"] == 1 and st["fallback_queries"] == 0


@pytest.mark.parametrize("storage,metric,engine", [("f32", 0, 0), ("f32", 1, 0), ("bf16", 0, 0), ("f32", 0, 1)])
def test_thousands_of_tied_rows_end_in_the_exhaustive_scan(oracle, storage, metric, engine):
    """3000 copies of one row around rank k: the range pass cannot hold every row that reaches its bound
    (buffers of 96 entries per scoring CTA, 2048 rows per query), so the query goes on to the exhaustive DOT64
    scan, which has no error bound and cannot fail.  The FMA engine (engine=1) has no range pass and hands
    over directly.  Result: still the lowest ids, bit-identical scores."""
    x, q = _data(oracle, 40_000, 128, 37, 61)
    x[5000:8000] = x[3]
    q[5] = x[3]
    q[20] = x[3]
    idx = b2r.IndexFlat(128, metric, storage=storage)
    idx.set_engine(engine)
    idx.add(x)
    D, I = idx.search(q, 10)
    xo, qo = (oracle.round_bf16(x), oracle.round_bf16(q)) if storage == "bf16" else (x, q)
    D0, I0 = oracle.search_ref(xo, qo, 10, metric, extra=4000)
    _check(D, I, D0, I0)
    assert list(I[5]) == [3] + list(range(5000, 5009)) and list(I[20]) == list(I[5])
    st = idx.last_stats()
    assert st["fallback_queries"] == 2 and st["exhaustive_queries"] == 2 and st["unverified"] == 0


def test_constant_corpus_and_large_k(oracle):
    """Every row identical (all scores tie): ids 0..k-1 for every query, through the exhaustive scan; k = 200."""
    x = np.tile(np.linspace(-1, 1, 96, dtype=np.float32), (5000, 1))
    q = np.random.default_rng(3).standard_normal((9, 96)).astype(np.float32)
    for k in (3, 200):
        idx = b2r.IndexFlatIP(96)
        idx.add(x)
        D, I = idx.search(q, k)
        assert np.array_equal(I, np.tile(np.arange(k), (9, 1)))
        D0, _ = oracle.search_defined(x[:k], q, k, 0)
        assert np.array_equal(_bits(D), _bits(D0))
        st = idx.last_stats()
        assert st["exhaustive_queries"] == 9 and st["unverified"] == 0


@pytest.mark.parametrize("storage,nq", [("f32", 300), ("bf16", 130), ("f32", 2100)])
def test_multi_round_scan_of_a_large_shard(oracle, monkeypatch, storage, nq):
    """Very large shards are scanned in rounds of growing size with the thresholds tightened in between (every
    CTA's buffers persist from round to round).  The knobs that decide it are lowered here so that a 300k-row
    corpus takes the same path as a 100M-row one: sampling launch, a first round of 256 tiles, tightening, the rest."""
    monkeypatch.setenv("B2R_SEED_TILES", "32")
    monkeypatch.setenv("B2R_ROUND_MIN", "1")
    x, q = _data(oracle, 300_000, 128, nq, 63)
    x[250_000:250_030] = x[9]                       # a tie cluster that only the last round sees
    q[1] = x[9]
    idx = b2r.IndexFlatIP(128, storage=storage)
    idx.add(x)
    D, I = idx.search(q, 10)
    xo, qo = (oracle.round_bf16(x), oracle.round_bf16(q)) if storage == "bf16" else (x, q)
    _check(D, I, *oracle.search_ref(xo, qo, 10, 0, extra=128))
    st = idx.last_stats()
    assert st["unverified"] == 0 and st["launches"] >= 10      # prep, sample, thr, 2 rounds + tightening, merge, hand-over chain


@pytest.mark.parametrize("nq,metric,k", [(800, b2r.METRIC_L2, 10), (512, b2r.METRIC_INNER_PRODUCT, 21), (2300, b2r.METRIC_INNER_PRODUCT, 5)])
def test_cta_pair_kernel_shapes(oracle, nq, metric, k):
    """fp32 rows with several query tiles run on the CTA-pair kernel (cta_group::2): an odd number of query tiles
    (800 queries = 7 tiles, the last pair half empty), the L2 metric (row norms shared by the pair), K' = 42, and a
    call of two launches whose second holds a single query tile."""
    x, q = _data(oracle, 70_000, 96, nq, 64)
    x[40_000:40_040] = x[5]                         # a tie cluster for query 3: hand-over from the pair kernel's buffers
    q[3] = x[5]
    idx = b2r.IndexFlat(96, metric)
    idx.add(x)
    D, I = idx.search(q, k)
    _check(D, I, *oracle.search_ref(x, q, k, metric, extra=128))
    st = idx.last_stats()
    assert st["engine"] == b2r.ENGINE_UMMA and st["unverified"] == 0 and st["fallback_queries"] >= 1


@pytest.mark.parametrize("storage", ["f32", "bf16"])
def test_every_query_of_a_large_batch_is_handed_over(oracle, storage):
    """2500 queries that all hit a cluster of 60 identical rows: none can be proven by the first pass, so the range
    list holds 2500 entries — more than one scoring launch takes (2048), i.e. two range launches and a merge that
    strides over both chunks of the candidate scratch."""
    x, _ = _data(oracle, 30_000, 64, 4, 65)
    x[2000:2060] = x[1]
    rng = np.random.default_rng(66)
    q = np.tile(x[1], (2500, 1)) + 1e-3 * rng.standard_normal((2500, 64)).astype(np.float32)
    q = oracle.normalize(q.astype(np.float32))
    idx = b2r.IndexFlatIP(64, storage=storage)
    idx.add(x)
    D, I = idx.search(q, 10)
    xo, qo = (oracle.round_bf16(x), oracle.round_bf16(q)) if storage == "bf16" else (x, q)
    _check(D, I, *oracle.search_ref(xo, qo, 10, 0, extra=128))
    st = idx.last_stats()
    assert st["fallback_queries"] == 2500 and st["exhaustive_queries"] == 0 and st["unverified"] == 0
    assert set(I[:, :10].ravel()) <= {1} | set(range(2000, 2060))


def test_second_range_pass_settles_what_the_first_cannot_hold(oracle):
    """3000 near-copies of one row at distances spread up to 0.08: the 1xTF32 range pass's band (4e-3) holds all of
    them and overflows, the 3xTF32 pass's band (1.4e-4) holds a few hundred — the query is settled there, without
    the exhaustive scan."""
    x, q = _data(oracle, 40_000, 128, 20, 67)
    rng = np.random.default_rng(68)
    noise = oracle.normalize(rng.standard_normal((3000, 128)).astype(np.float32))
    radius = rng.uniform(0.0, 0.08, (3000, 1)).astype(np.float32)
    x[5000:8000] = oracle.normalize((x[3][None, :] + radius * noise).astype(np.float32))
    q[7] = x[3]
    idx = b2r.IndexFlatIP(128)
    idx.add(x)
    D, I = idx.search(q, 10)
    _check(D, I, *oracle.search_ref(x, q, 10, 0, extra=4000))
    st = idx.last_stats()
    assert st["fallback_queries"] >= 1 and st["exhaustive_queries"] == 0 and st["unverified"] == 0


def test_fallback_counters_add_up_over_query_batches(oracle):
    """A call of more than 8192 queries is walked in batches; an unproven query in the second batch is handled
    there and the counters of the batches add up on the device (no read-back between them)."""
    x, q = _data(oracle, 9000, 64, 9000, 62)
    x[100:150] = x[2]
    q[8500] = x[2]
    q[17] = x[2]
    idx = b2r.IndexFlatIP(64)
    idx.add(x)
    D, I = idx.search(q, 5)
    _check(D, I, *oracle.search_ref(x, q, 5, 0, extra=128))
    st = idx.last_stats()
    assert st["nq"] == 9000 and st["fallback_queries"] >= 2 and st["unverified"] == 0


@pytest.mark.parametrize("metric", [b2r.METRIC_INNER_PRODUCT, b2r.METRIC_L2])
def test_3xtf32_engine_parity(oracle, metric):
    """The split-TF32 variant of the tcgen05 kernel (exact mode's fallback), forced, against the oracle."""
    x, q = _data(oracle, 50_000, 768, 150, 70)
    idx = b2r.IndexFlat(768, metric)
    idx.set_engine(b2r.ENGINE_UMMA3)
    idx.add(x)
    D, I = idx.search(q, 10)
    _check(D, I, *oracle.search_ref(x, q, 10, metric))
    st = idx.last_stats()
    assert st["engine"] == b2r.ENGINE_UMMA3 and st["unverified"] == 0 and st["kprime"] == 16
    fast = b2r.IndexFlatIP(768, storage="bf16")
    fast.set_engine(b2r.ENGINE_UMMA3)
    fast.add(x[:100])
    with pytest.raises(b2r.B2RError):                 # 3xTF32 is for fp32 rows
        fast.search(q, 3)


def test_full_size_properties_1m_768(oracle):
    """configs[1] at full size (1M x 768): properties that need no full oracle run."""
    import torch
    n, d, k = 1_000_000, 768, 10
    g = torch.Generator(device="cuda")
    g.manual_seed(123)
    x = torch.randn((n, d), generator=g, device="cuda")
    b2r.normalize_L2(x)
    sel = torch.arange(0, n, n // 96, device="cuda")[:96]
    q = x[sel].contiguous()
    exact = b2r.IndexFlatIP(d)
    exact.add(x)
    D, I = exact.search(q, k)
    D, I = D.cpu().numpy(), I.cpu().numpy()
    assert np.array_equal(I[:, 0], sel.cpu().numpy())                    # a stored row retrieves itself first
    assert np.all(np.abs(D[:, 0] - 1.0) < 1e-6)
    assert np.all(D[:, :-1] >= D[:, 1:])                                 # best first
    assert all(len(set(r)) == k for r in I)                             # no id twice
    assert exact.last_stats()["unverified"] == 0
    # the whole batch against the defined oracle on the same rows (96 queries x 1M rows, fp64 BLAS + DOT64)
    xh = x.cpu().numpy()
    D0, I0 = oracle.search_ref(xh, q.cpu().numpy(), k, 0)
    _check(D, I, D0, I0)
    # idempotence: same call, same bits
    D2, I2 = exact.search(q, k)
    assert np.array_equal(I2.cpu().numpy(), I) and np.array_equal(_bits(D2.cpu().numpy()), _bits(D))
    # row-sharded (G=4) == unsharded, through the packed merge
    packs = []
    for r in range(4):
        lo, hi = b2r.shard_bounds(n, 4, r)
        sh = b2r.IndexFlatIP(d)
        sh.add(x[lo:hi])
        packs.append(sh.search_packed(q, k, lo))
        del sh
    Dm, Im = b2r.merge_packed(torch.stack(packs), 0)
    assert np.array_equal(Im.cpu().numpy(), I) and np.array_equal(_bits(Dm.cpu().numpy()), _bits(D))
    # fast mode: recall against exact mode
    fast = b2r.IndexFlatIP(d, storage="bf16")
    fast.add(x)
    _, If = fast.search(q, k)
    rec = np.mean([len(set(a) & set(b)) / k for a, b in zip(If.cpu().numpy(), I)])
    assert rec >= 0.9


@pytest.mark.parametrize("n,d,nq,k", [(1, 8, 1, 1), (33, 100, 3, 7), (300, 64, 130, 10), (4097, 260, 17, 33)])
def test_bf16_ragged_small_shapes(oracle, n, d, nq, k):
    """Fast mode on the tcgen05 engine with fewer rows than a tile, ragged d, more than one query tile."""
    rng = np.random.default_rng(n + d)
    x = rng.standard_normal((n, d)).astype(np.float32)
    q = rng.standard_normal((nq, d)).astype(np.float32)
    for metric in (b2r.METRIC_INNER_PRODUCT, b2r.METRIC_L2):
        idx = b2r.IndexFlat(d, metric, storage="bf16")
        idx.add(x)
        D, I = idx.search(q, k)
        _check(D, I, *oracle.search_defined(oracle.round_bf16(x), oracle.round_bf16(q), k, metric))
        assert idx.last_stats()["engine"] == b2r.ENGINE_UMMA


def test_maximum_k(oracle):
    """k = 1024 (the documented maximum) goes through the FMA engine; k = 128 is the largest exact-mode k on tcgen05."""
    x, q = _data(oracle, 6000, 128, 5, 80)
    idx = b2r.IndexFlatIP(128)
    idx.add(x)
    D, I = idx.search(q, 1024)
    _check(D, I, *oracle.search_defined(x, q, 1024, 0))
    assert idx.last_stats()["engine"] == b2r.ENGINE_FMA
    D, I = idx.search(q, 128)
    _check(D, I, *oracle.search_defined(x, q, 128, 0))
    assert idx.last_stats()["engine"] in (b2r.ENGINE_UMMA, b2r.ENGINE_UMMA3)
    with pytest.raises(b2r.B2RError):
        idx.search(q, 1025)


@pytest.mark.parametrize("metric", [b2r.METRIC_INNER_PRODUCT, b2r.METRIC_L2])
def test_exact_mode_with_bf16_shadow(oracle, metric):
    """Exact mode scoring a bf16 shadow copy: same bits as plain exact mode (fp32 rows are re-scored)."""
    x, q = _data(oracle, 80_000, 768, 40, 90)
    x[5000:5040] = x[11]                       # a tie cluster: forces the proof to fail for query 0
    q[0] = x[11]
    idx = b2r.IndexFlat(768, metric, shadow=True)
    idx.add(x[:50_000])
    idx.add(x[50_000:])                        # growth carries the shadow along
    D, I = idx.search(q[1:], 10)
    _check(D, I, *oracle.search_ref(x, q[1:], 10, metric))
    st = idx.last_stats()
    assert st["engine"] == b2r.ENGINE_SHADOW and st["unverified"] == 0 and st["fallback"] == 0
    D, I = idx.search(q, 10)                    # with the tie cluster: query 0 takes the range pass, still exact
    _check(D, I, *oracle.search_ref(x, q, 10, metric, extra=128))
    st = idx.last_stats()
    assert st["fallback"] == 1 and st["fallback_queries"] == 1 and st["unverified"] == 0
    assert np.array_equal(idx.reconstruct(77), x[77])
    with pytest.raises(b2r.B2RError):
        b2r.IndexFlatIP(64, storage="bf16", shadow=True)


def test_random_shapes_against_the_oracle(oracle):
    """40 seeded random (n, d, nq, k, metric, storage, engine) cases, small enough for the exhaustive oracle."""
    rng = np.random.default_rng(2026)
    for case in range(40):
        n = int(rng.integers(1, 3000))
        d = int(rng.integers(1, 300))
        nq = int(rng.integers(1, 300))
        k = int(rng.integers(1, 50))
        metric = int(rng.integers(0, 2))
        storage = ["f32", "bf16"][int(rng.integers(0, 2))]
        engine = [b2r.ENGINE_AUTO, b2r.ENGINE_FMA, b2r.ENGINE_UMMA3][int(rng.integers(0, 3))]
        if storage == "bf16" and engine == b2r.ENGINE_UMMA3:
            engine = b2r.ENGINE_AUTO
        x = rng.standard_normal((n, d)).astype(np.float32)
        q = rng.standard_normal((nq, d)).astype(np.float32)
        if case % 3 == 0 and n > 10:
            x[n // 2:] = x[: n - n // 2]                 # duplicate rows: exact ties
        idx = b2r.IndexFlat(d, metric, storage=storage, shadow=(storage == "f32" and case % 4 == 1))
        idx.set_engine(engine)
        idx.add(x)
        D, I = idx.search(q, k)
        xo, qo = (oracle.round_bf16(x), oracle.round_bf16(q)) if storage == "bf16" else (x, q)
        D0, I0 = oracle.search_defined(xo, qo, k, metric)
        assert np.array_equal(I, I0), (case, n, d, nq, k, metric, storage, engine)
        assert np.array_equal(_bits(D), _bits(D0)), (case, n, d, nq, k, metric, storage, engine)


def test_more_queries_than_one_batch(oracle):
    """20k queries in one call: the C ABI walks them in batches of 8192 with a bounded scratch."""
    x, q = _data(oracle, 6000, 64, 20_000, 95)
    for storage in ("f32", "bf16"):
        idx = b2r.IndexFlatIP(64, storage=storage)
        idx.add(x)
        D, I = idx.search(q, 5)
        xo, qo = (oracle.round_bf16(x), oracle.round_bf16(q)) if storage == "bf16" else (x, q)
        _check(D, I, *oracle.search_ref(xo, qo, 5, 0))
        st = idx.last_stats()
        assert st["nq"] == 20_000 and st["unverified"] == 0


def test_packed_shard_search_is_asynchronous_and_exact(oracle):
    """The sharded path's payload: packed shard-local hits; an unprovable query in one shard is settled on the
    device before the payload is read (no flag travels, no host check)."""
    import torch
    x, q = _data(oracle, 40_001, 128, 33, 97)
    x[300:340] = x[5]                                   # tie cluster: shard 0's first pass cannot prove query 0
    q[0] = x[5]
    D0, I0 = oracle.search_ref(x, q, 10, 0, extra=128)
    for G in (2, 3):
        bufs, fb = [], 0
        for r in range(G):
            lo, hi = b2r.shard_bounds(len(x), G, r)
            sh = b2r.IndexFlatIP(128)
            sh.add(x[lo:hi])
            bufs.append(sh.search_packed(q, 10, lo))
            fb += sh.last_stats()["fallback_queries"]
        D, I = b2r.merge_packed(torch.stack(bufs), 0)
        _check(D.cpu().numpy(), I.cpu().numpy(), D0, I0)
        assert fb >= 1
        padded = torch.stack([torch.cat([b.flatten(), b.new_zeros(3)]) for b in bufs])      # lists 333 words apart
        D, I = b2r.merge_packed_strided(padded, 33, 10, 0)
        _check(D.cpu().numpy(), I.cpu().numpy(), D0, I0)

"""CPU tests: the oracle against the reference-generated golden vectors and its own known answers.

The reference holds no test for the search path; the golden vectors in
tests/golden/reward_cosine.npz come from the reference's own RewardModel methods (see
tests/golden/make_reward_golden.py) and pin the normalise -> inner product -> max arithmetic.
"""
import os

import numpy as np
import pytest

FLT_MAX = float(np.finfo(np.float32).max)


def _ulp_diff(a, b):
    a = np.asarray(a, dtype=np.float32).view(np.int32).astype(np.int64)
    b = np.asarray(b, dtype=np.float32).view(np.int32).astype(np.int64)
    a = np.where(a < 0, -(a & 0x7fffffff), a)
    b = np.where(b < 0, -(b & 0x7fffffff), b)
    return np.abs(a - b)


def test_dot64_c_matches_numpy_restatement(oracle):
    rng = np.random.default_rng(0)
    for d in (1, 3, 4, 5, 31, 100, 128, 129, 384, 768):
        a = rng.standard_normal(d).astype(np.float32)
        b = rng.standard_normal(d).astype(np.float32)
        assert oracle.dot64(a, b) == oracle.dot64_numpy(a, b)
        assert abs(oracle.dot64(a, b) - float(a.astype(np.float64) @ b.astype(np.float64))) < 1e-12 * d


def test_normalize_matches_sklearn(oracle):
    from sklearn.preprocessing import normalize as sk_normalize
    rng = np.random.default_rng(1)
    x = rng.standard_normal((64, 768)).astype(np.float32) * rng.uniform(1e-3, 1e3, (64, 1)).astype(np.float32)
    x[5] = 0.0
    y = oracle.normalize(x)
    ref = sk_normalize(x)
    assert np.all(y[5] == 0.0) and np.all(ref[5] == 0.0)          # zero-norm row unchanged (reference :70 via sklearn)
    assert _ulp_diff(y, ref).max() <= 8                            # sklearn accumulates the norm in fp32
    x64 = x.astype(np.float64)
    nz = np.linalg.norm(x64, axis=1, keepdims=True)
    nz[nz == 0] = 1
    assert _ulp_diff(y, (x64 / nz).astype(np.float32)).max() <= 1  # vs the exactly rounded result
    n = np.sqrt((y.astype(np.float64) ** 2).sum(1))
    assert np.all(np.abs(np.delete(n, 5) - 1) < 1e-6)


def test_round_bf16_matches_torch(oracle):
    import torch
    rng = np.random.default_rng(2)
    x = np.concatenate([rng.standard_normal(4096), [0.0, -0.0, 1.0, 3.3895314e38, 1e-40, np.inf]]).astype(np.float32)
    ref = torch.from_numpy(x).to(torch.bfloat16).to(torch.float32).numpy()
    assert np.array_equal(oracle.round_bf16(x).view(np.uint32), ref.view(np.uint32))


def test_reward_golden_from_reference(oracle, golden_dir):
    """Oracle cosine/max vs values produced by the reference's RewardModel (sklearn cosine_similarity)."""
    g = np.load(os.path.join(golden_dir, "reward_cosine.npz"))
    for ci in range(int(g["ncases"])):
        resp, docs, query, gt = (g[f"c{ci}_{k}"] for k in ("resp", "docs", "query", "gt"))
        cos, mx = oracle.cosine_rows(resp, docs)
        np.testing.assert_allclose(mx, g[f"c{ci}_fact"], rtol=0, atol=2e-6)      # reference :63-71
        _, rel = oracle.cosine_rows(resp, query[:, None, :])
        np.testing.assert_allclose(rel, g[f"c{ci}_rel"], rtol=0, atol=2e-6)      # reference :73-79
        _, gts = oracle.cosine_rows(resp, gt[:, None, :])
        np.testing.assert_allclose(gts, g[f"c{ci}_gtsim"], rtol=0, atol=2e-6)    # reference :99-104
        # the same numbers through the search path: normalise, IndexFlatIP, k=1
        for i in range(resp.shape[0]):
            if docs.shape[1] == 0:
                continue
            xs, qs = oracle.normalize(docs[i]), oracle.normalize(resp[i:i + 1])
            D, I = oracle.search_defined(xs, qs, 1, oracle.METRIC_IP)
            assert abs(float(D[0, 0]) - float(g[f"c{ci}_fact"][i])) <= 2e-6
            assert I[0, 0] == int(np.argmax(cos[i])) or abs(cos[i, I[0, 0]] - cos[i].max()) <= 2e-6


def test_search_golden_regression(oracle, golden_dir):
    g = np.load(os.path.join(golden_dir, "search_small.npz"))
    for ci in range(int(g["ncases"])):
        k, metric = (int(v) for v in g[f"c{ci}_meta"])
        D, I = oracle.search_defined(g[f"c{ci}_x"], g[f"c{ci}_q"], k, metric)
        assert np.array_equal(I, g[f"c{ci}_I"])
        assert np.array_equal(D.view(np.uint32), g[f"c{ci}_D"].view(np.uint32))
        D2, I2 = oracle.search_ref(g[f"c{ci}_x"], g[f"c{ci}_q"], k, metric)
        assert np.array_equal(I2, I) and np.array_equal(D2.view(np.uint32), D.view(np.uint32))


def test_identity_corpus(oracle):
    d = 48
    x = np.eye(d, dtype=np.float32)
    rng = np.random.default_rng(3)
    q = rng.standard_normal((5, d)).astype(np.float32)
    D, I = oracle.search_defined(x, q, d, oracle.METRIC_IP)
    for i in range(5):
        assert np.array_equal(I[i], np.argsort(-q[i], kind="stable"))
        assert np.array_equal(D[i], q[i][I[i]])


def test_ties_lower_id_first_and_padding(oracle):
    rng = np.random.default_rng(4)
    base = rng.standard_normal((10, 32)).astype(np.float32)
    x = np.concatenate([base, base, base])            # ids i, i+10, i+20 tie exactly
    q = base[:3].copy()
    for metric in (oracle.METRIC_IP, oracle.METRIC_L2):
        D, I = oracle.search_defined(x, q, 6, metric)
        for r in range(3):
            ids = I[r]
            for a in range(0, 6, 3):
                assert ids[a] + 10 == ids[a + 1] and ids[a] + 20 == ids[a + 2]
                assert D[r, a] == D[r, a + 1] == D[r, a + 2]
    D, I = oracle.search_defined(x[:4], q, 7, oracle.METRIC_IP)
    assert np.all(I[:, 4:] == -1) and np.all(D[:, 4:] == -FLT_MAX)
    D, I = oracle.search_defined(x[:4], q, 7, oracle.METRIC_L2)
    assert np.all(I[:, 4:] == -1) and np.all(D[:, 4:] == FLT_MAX)
    D, I = oracle.search_ref(x[:0], q, 3, oracle.METRIC_IP)
    assert np.all(I == -1)


def test_ip_l2_rank_equivalence_on_unit_vectors(oracle):
    rng = np.random.default_rng(5)
    x = oracle.normalize(rng.standard_normal((2000, 64)).astype(np.float32))
    q = oracle.normalize(rng.standard_normal((8, 64)).astype(np.float32))
    Di, Ii = oracle.search_defined(x, q, 10, oracle.METRIC_IP)
    Dl, Il = oracle.search_defined(x, q, 10, oracle.METRIC_L2)
    assert np.array_equal(Ii, Il)
    np.testing.assert_allclose(Dl, 2 - 2 * Di, atol=2e-6)


def test_permutation_property(oracle):
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=15, deadline=None)
    @given(st.integers(0, 2 ** 31 - 1))
    def run(seed):
        rng = np.random.default_rng(seed)
        n, d, k = 200, 24, 7
        x = rng.standard_normal((n, d)).astype(np.float32)
        q = rng.standard_normal((3, d)).astype(np.float32)
        perm = rng.permutation(n)
        D0, I0 = oracle.search_defined(x, q, k)
        D1, I1 = oracle.search_defined(x[perm], q, k)
        assert np.array_equal(D0, D1)
        assert np.array_equal(perm[I1], I0)          # no exact ties in random data
    run()


def test_sharded_equals_unsharded_on_oracle(oracle):
    rng = np.random.default_rng(6)
    n, d, k = 3001, 40, 9
    x = rng.standard_normal((n, d)).astype(np.float32)
    x[1500:1510] = x[10:20]                           # ties across shards
    q = np.concatenate([rng.standard_normal((4, d)).astype(np.float32), x[10:12]])
    D0, I0 = oracle.search_defined(x, q, k)
    for G in (2, 4, 8):
        per = -(-n // G)
        Ds, Is = [], []
        for r in range(G):
            lo, hi = min(n, r * per), min(n, (r + 1) * per)
            D, I = oracle.search_defined(x[lo:hi], q, k)
            Ds.append(D)
            Is.append(np.where(I >= 0, I + lo, -1))
        Dc, Ic = np.concatenate(Ds, 1), np.concatenate(Is, 1)
        order = np.lexsort((Ic, -Dc.astype(np.float64)), axis=1)
        Dm = np.take_along_axis(Dc, order, 1)[:, :k]
        Im = np.take_along_axis(Ic, order, 1)[:, :k]
        assert np.array_equal(Im, I0) and np.array_equal(Dm, D0)


def test_fp32_blas_baseline_agrees(oracle):
    rng = np.random.default_rng(8)
    x = oracle.normalize(rng.standard_normal((20000, 128)).astype(np.float32))
    q = oracle.normalize(rng.standard_normal((33, 128)).astype(np.float32))
    for metric in (oracle.METRIC_IP, oracle.METRIC_L2):
        D0, I0, gaps = oracle.search_ref(x, q, 10, metric, return_gaps=True)
        D1, I1 = oracle.search_fp32_blas(x, q, 10, metric)
        clear = gaps > 1e-5                       # fp32 engines may legitimately flip nearer ties
        assert clear.sum() >= 30
        assert np.array_equal(I0[clear], I1[clear])
        np.testing.assert_allclose(D1, D0, rtol=1e-5, atol=2e-6)


def test_synth_rows_chunk_reproducible(oracle):
    a = oracle.synth_rows(0, 65000, 2000, 16)
    b = np.concatenate([oracle.synth_rows(0, 65000, 700, 16), oracle.synth_rows(0, 65700, 1300, 16)])
    assert np.array_equal(a, b)
    assert not np.array_equal(a, oracle.synth_rows(1, 65000, 2000, 16))

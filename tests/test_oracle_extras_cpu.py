"""CPU checks of the oracle's helpers that the full-size parity tests and the timed CPU baseline rely on:
the chunked fp64 accumulator (oracle_acc_*), the cache-blocked fp32 batch search and the IVF_PQ search over a
given partition list — each against the plain single-query oracle (oracle_flat_search / oracle_ivfpq_search),
which is the function pinned by the reference's golden vectors (tests/test_oracle_golden.py)."""
import numpy as np
import pytest

from parity import assert_topk_parity


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_chunked_accumulator_equals_flat_search(oracle_mod, metric):
    rng = np.random.default_rng(3)
    n, d, nq, k = 20_000, 48, 9, 17
    X = rng.standard_normal((n, d), dtype=np.float32)
    X[5000] = X[4000]  # an exact duplicate: the tie rule (ascending label) must survive chunking
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    Q[0] = X[4000]
    allow = oracle_mod.make_bitmap(rng.random(n) < 0.4)
    dele = oracle_mod.make_bitmap(rng.random(n) < 0.2)
    for use_allow, use_del in ((False, False), (True, True)):
        acc = oracle_mod.FlatAccumulator(Q, k, metric, label_base=500)
        for r0 in range(0, n, 3333):  # ragged chunks
            acc.add(X[r0:r0 + 3333], row0=r0, allow=allow if use_allow else None, deleted=dele if use_del else None)
        res = acc.finish()
        for j in range(nq):
            wl, _, wd64 = oracle_mod.flat_search(X, Q[j], k, metric, allow if use_allow else None,
                                                 dele if use_del else None, label_base=500)
            assert list(res[j][0]) == list(wl)
            np.testing.assert_allclose(res[j][1], wd64, rtol=1e-12, atol=1e-14)


def test_accumulator_allow_bits_cuts_rows_beyond_the_bitmap(oracle_mod):
    rng = np.random.default_rng(4)
    n, d = 1000, 8
    X = rng.standard_normal((n, d), dtype=np.float32)
    q = rng.standard_normal(d, dtype=np.float32)
    allow = oracle_mod.make_bitmap(np.ones(640, bool))
    acc = oracle_mod.FlatAccumulator(q, n, "l2")
    acc.add(X, allow=allow, allow_bits=600)
    (labels, _), = acc.finish()
    assert len(labels) == 600 and labels.max() == 599


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
@pytest.mark.parametrize("nq", [1, 16, 37])
def test_blocked_batch_search_matches_truth(oracle_mod, metric, nq):
    rng = np.random.default_rng(5 + nq)
    n, d, k = 9_001, 100, 10
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    allow = oracle_mod.make_bitmap(rng.random(n) < 0.5)
    L, D, C = oracle_mod.flat_search_batch_blocked(X, Q, k, metric, allow=allow, nthreads=3)
    for j in range(nq):
        wl, _, wd64 = oracle_mod.flat_search(X, Q[j], k, metric, allow)
        assert C[j] == len(wl)
        truth = lambda lab: oracle_mod.distance_f64(metric, Q[j], X[lab])
        # |q|^2 + |x|^2 - 2 q.x carries ~1e-6 of (|q|^2 + |x|^2): ids may swap inside that band
        got_l, got_d = L[j, :C[j]], D[j, :C[j]]
        for g, w in zip(got_l, wl):
            if g != w:
                assert abs(truth(int(g)) - truth(int(w))) <= 1e-4 * max(1.0, abs(truth(int(w))))
        np.testing.assert_allclose(got_d, wd64, rtol=1e-4, atol=2e-4)


def test_blocked_batch_search_edge_cases(oracle_mod):
    rng = np.random.default_rng(6)
    X = rng.standard_normal((5, 4), dtype=np.float32)
    Q = rng.standard_normal((3, 4), dtype=np.float32)
    L, D, C = oracle_mod.flat_search_batch_blocked(X, Q, 10, "l2")  # k > n
    assert (C == 5).all()
    for j in range(3):
        wl, wd, _ = oracle_mod.flat_search(X, Q[j], 10, "l2")
        assert list(L[j, :5]) == list(wl)
    L, D, C = oracle_mod.flat_search_batch_blocked(X, Q, 2, "dot", allow=oracle_mod.make_bitmap(np.zeros(5, bool)))
    assert (C == 0).all()


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_ivfpq_search_over_given_partitions(oracle_mod, metric):
    rng = np.random.default_rng(8)
    n, d, P, M = 3000, 32, 12, 8
    X = rng.standard_normal((n, d), dtype=np.float32)
    ix = oracle_mod.IvfPqIndex.train(X, metric, P, M, iters=3)
    q = rng.standard_normal(d, dtype=np.float32)
    want = ix.search(X, q, 10, 5, 2)
    got = ix.search_parts(X, q, 10, want["parts"], 2)  # the probe's own list: identical everything
    for key in ("cand_labels", "cand_adc", "labels", "dist"):
        np.testing.assert_array_equal(got[key], want[key])
    # order of the list does not matter; refine_factor 0 reads no row at all (X may be omitted)
    g0 = ix.search_parts(None, q, 10, want["parts"][::-1].copy(), 0)
    w0 = ix.search(X, q, 10, 5, 0)
    np.testing.assert_array_equal(g0["labels"], w0["labels"])
    np.testing.assert_array_equal(g0["dist"], w0["dist"])
    # a different list gives the candidates of those lists only
    other = np.array([p for p in range(P) if p not in set(want["parts"].tolist())][:3], np.int32)
    g1 = ix.search_parts(X, q, 10, other, 1)
    members = set()
    for p in other:
        members |= set(ix.row_ids[ix.part_offsets[p]:ix.part_offsets[p + 1]].tolist())
    assert set(g1["labels"].tolist()) <= members

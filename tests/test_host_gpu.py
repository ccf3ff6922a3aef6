"""The reference's SQL tests replayed through the C++ host mirror (csrc/host/): LanceIndex::Search
(src/lance_index.cpp:442-465), the lance_search table function (src/lance_search.cpp:33-104) and the optimizer's
replacement scan with predicate push-down (src/lance_optimizer.cpp:35-109, 204-344) — on the GPU, via the C ABI."""
import json
import os

import numpy as np
import pytest

from parity import assert_topk_parity

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
FX = json.load(open(os.path.join(HERE, "golden", "reference_sql_fixtures.json")))


def build_index(host, case, rowid_base=0):
    X = np.array(case["rows"], np.float32)
    ix = host.HostIndex(X.shape[1], case["metric"])
    rowids = np.arange(len(X), dtype=np.int64) + rowid_base
    ix.append(X, rowids)
    for name, vals in case.get("columns", {}).items():
        ix.columns.append(name, vals)
    if "deleted_ids" in case:
        ids = np.array(case["ids"])
        ix.delete(rowids[np.isin(ids, case["deleted_ids"])])
        assert ix.count() == len(X) - len(case["deleted_ids"])
    return ix, X, rowids


@pytest.mark.parametrize("case", FX["cases"], ids=[c["name"] for c in FX["cases"]])
def test_reference_sql_fixture_through_host_mirror(gpu, host, case):
    ix, X, rowids = build_index(host, case)
    ids = np.array(case["ids"])
    q = np.array(case["query"], np.float32)
    with ix:
        if "predicate" in case or "lance_optimizer" in case["name"]:
            # SELECT id FROM t [WHERE p] ORDER BY array_distance(vec, q) LIMIT k, rewritten to lance_index_scan
            got_rows = ix.lance_index_scan(q, case["k"], case.get("predicate", ""))
            _, got_d = ix.search(q, case["k"], case.get("predicate", ""))
        else:
            # SELECT row_id, distance FROM lance_search('t', 'idx', q, k)
            got_rows, got_d, chunks = ix.lance_search(q, case["k"])
            assert chunks == (1 if len(got_rows) else 0)
        if "expect_ids" in case:
            # position by position, except between rows whose distances tie (ids 3 and 4 both sit at 2.0 in
            # "filter score > 20": the reference's own order there is whatever its heap emitted)
            Xd, qd = X.astype(np.float64), q.astype(np.float64)
            truth = lambda lab: float(((Xd[lab] - qd) ** 2).sum())
            want = [list(ids).index(i) for i in case["expect_ids"]]
            assert_topk_parity(got_rows, got_d, want, [truth(l) for l in want], truth, what=case["name"])
        if "expect_row_ids" in case:
            assert list(got_rows) == case["expect_row_ids"]
        if "expect_count" in case:
            assert len(got_rows) == case["expect_count"]
        if "expect_distances" in case:
            assert ["%.6f" % v for v in got_d] == ["%.6f" % v for v in case["expect_distances"]]
        if "expect_ids_sorted_by_distance_then_any" in case:
            pos = 0
            for group in case["expect_ids_sorted_by_distance_then_any"]:
                assert sorted(ids[got_rows[pos:pos + len(group)]]) == sorted(group)
                pos += len(group)


def test_wrong_dimension_returns_no_rows(gpu, host):
    """test/sql/lance_basic.test:55-59: a query of the wrong dimension yields 0 rows, not an error."""
    with host.HostIndex(3, "l2") as ix:
        ix.append(np.eye(3, dtype=np.float32), [0, 1, 2])
        rows, d, chunks = ix.lance_search([1.0, 0.0], 1)
        assert len(rows) == 0 and chunks == 0
        assert len(ix.lance_index_scan([1.0, 0.0, 0.0, 0.0], 2)) == 0


def test_rowid_mapping_append_delete_and_chunked_scan(gpu, host, oracle_mod):
    """Row ids are DuckDB's, not labels: non-contiguous row ids, delete by row id, rows appended later, and a
    k > 2048 result emitted in STANDARD_VECTOR_SIZE chunks (src/lance_search.cpp:84-104)."""
    rng = np.random.default_rng(21)
    n, d = 6000, 24
    X = rng.standard_normal((n, d), dtype=np.float32)
    rowids = rng.permutation(10 * n)[:n].astype(np.int64)
    q = rng.standard_normal(d, dtype=np.float32)
    with host.HostIndex(d, "l2") as ix:
        ix.append(X[:4000], rowids[:4000])
        ix.append(X[4000:], rowids[4000:])  # appended after "index creation" (lance_basic.test:45-53)
        dele = rng.choice(n, 300, replace=False)
        ix.delete(rowids[dele])
        ix.delete(np.array([10 ** 9]))  # unknown row id: ignored
        assert ix.count() == n - 300
        deleted = np.zeros(n, bool)
        deleted[dele] = True
        k = 3000
        rows, dist, chunks = ix.lance_search(q, k)
        assert len(rows) == k and chunks == 2
        wl, wd32, wd64 = oracle_mod.flat_search(X, q, k, "l2", None, oracle_mod.make_bitmap(deleted))
        label_of = {int(r): i for i, r in enumerate(rowids)}
        got_labels = [label_of[int(r)] for r in rows]
        assert_topk_parity(got_labels, dist, wl, wd64, lambda lab: oracle_mod.distance_f64("l2", q, X[lab]), what="host chunked")
        assert (np.diff(dist) >= 0).all()  # result order = ascending distance (the ORDER BY node is gone)


def test_host_index_options_batch_and_ann(gpu, host, oracle_mod):
    """WITH (metric='cosine', nprobes=.., refine_factor=..) and lance_create_ann_index through the mirror."""
    rng = np.random.default_rng(23)
    n, d, k = 20000, 32, 10
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((12, d), dtype=np.float32)
    score = rng.integers(0, 100, n)
    with host.HostIndex(d, "cosine", nprobes=64, refine_factor=4) as ix:
        ix.append(X, np.arange(n))
        ix.columns.append("score", score.tolist())
        ids, dist, cnt = ix.search_batch(Q, k, "score > 20")
        allow = oracle_mod.make_bitmap(score > 20)
        for j in range(len(Q)):
            wl, wd32, wd64 = oracle_mod.flat_search(X, Q[j], k, "cosine", allow, None)
            assert_topk_parity(ids[j, :cnt[j]], dist[j, :cnt[j]], wl, wd64,
                               lambda lab: oracle_mod.distance_f64("cosine", Q[j], X[lab]), what=f"host batch q{j}")
        ix.create_ann_index(64, 8)  # SELECT * FROM lance_create_ann_index('t', 'idx', 64, 8)
        rows, dist1 = ix.search(Q[0], k)
        assert len(rows) == k and (np.diff(dist1) >= 0).all()
        # refine_factor >= 1: the returned distances are exact distances of the returned rows
        for r, dd in zip(rows, dist1):
            assert abs(oracle_mod.distance_f64("cosine", Q[0], X[r]) - dd) <= 1e-4 * max(dd, 1e-3)
        wl, wd32, wd64 = oracle_mod.flat_search(X, Q[0], k, "cosine", None, None)
        assert len(set(rows.tolist()) & set(wl.tolist())) >= 3  # coarse PQ (8 x 4-d sub-vectors): recall, not parity


def test_predicate_bitmap_cache_follows_appends(gpu, host, oracle_mod):
    """LanceIndex keeps predicate -> bitmap while labels and columns are unchanged; an append must invalidate it."""
    rng = np.random.default_rng(29)
    n, d, k = 3000, 16, 5
    X = rng.standard_normal((2 * n, d), dtype=np.float32)
    q = rng.standard_normal(d, dtype=np.float32)
    score = rng.integers(0, 100, 2 * n)
    with host.HostIndex(d, "l2") as ix:
        ix.append(X[:n], np.arange(n))
        ix.columns.append("score", score[:n].tolist())
        for _ in range(2):  # second call is served from the cache
            rows, dist = ix.search(q, k, "score < 30")
            wl, _, wd64 = oracle_mod.flat_search(X[:n], q, k, "l2", oracle_mod.make_bitmap(score[:n] < 30), None)
            assert list(rows) == list(wl)
        # a very near neighbour that passes the predicate arrives with the next append
        X[n + 5] = q + 1e-3
        score[n + 5] = 7
        ix.append(X[n:], np.arange(n, 2 * n))
        ix.columns.append("score", score[n:].tolist())
        rows, dist = ix.search(q, k, "score < 30")
        wl, _, wd64 = oracle_mod.flat_search(X, q, k, "l2", oracle_mod.make_bitmap(score < 30), None)
        assert rows[0] == n + 5 and list(rows) == list(wl)
        rows2, _ = ix.search(q, k, "score >= 30")  # another predicate, same index
        assert n + 5 not in rows2.tolist()


def test_merge_indexes_moves_live_rows_and_their_columns(gpu, host, oracle_mod):
    """LanceIndex::MergeIndexes (src/lance_index.cpp:624-668): the live rows of the other index land in this one
    (device-side lance_detached_merge), keep their DuckDB row ids and their extra column values; rows deleted from
    the other index stay out.  Searching afterwards equals the oracle over the union."""
    rng = np.random.default_rng(41)
    na, nb, d, k = 900, 600, 16, 6
    A = rng.standard_normal((na, d), dtype=np.float32)
    B = rng.standard_normal((nb, d), dtype=np.float32)
    sa = rng.integers(0, 100, na)
    sb = rng.integers(0, 100, nb)
    rows_a = np.arange(na)                 # DuckDB row ids of A
    rows_b = 10_000 + np.arange(nb)        # ... and of B (disjoint)
    dead_b = np.array([10_003, 10_250, 10_599])
    with host.HostIndex(d, "l2") as ia, host.HostIndex(d, "l2") as ib:
        ia.append(A, rows_a)
        ia.columns.append("score", sa.tolist())
        ib.append(B, rows_b)
        ib.columns.append("score", sb.tolist())
        ib.delete(dead_b)
        assert ia.merge_indexes(ib) == nb - 3
        assert ia.count() == na + nb - 3
        keep = ~np.isin(rows_b, dead_b)
        X = np.concatenate([A, B[keep]])
        rowids = np.concatenate([rows_a, rows_b[keep]])
        score = np.concatenate([sa, sb[keep]])
        for q in (B[250], B[7], A[11], rng.standard_normal(d).astype(np.float32)):
            rows, dist = ia.search(q, k)
            wl, _, wd64 = oracle_mod.flat_search(X, q, k, "l2")
            assert list(rows) == rowids[wl].tolist()
            np.testing.assert_allclose(dist, wd64, rtol=1e-4, atol=1e-6)
            rows, dist = ia.search(q, k, "score < 40")
            wl, _, wd64 = oracle_mod.flat_search(X, q, k, "l2", oracle_mod.make_bitmap(score < 40), None)
            assert list(rows) == rowids[wl].tolist()
        assert 10_250 not in ia.search(B[250], k)[0].tolist()  # deleted before the merge
        # merging an index without live rows is a no-op
        with host.HostIndex(d, "l2") as ie:
            assert ia.merge_indexes(ie) == 0


def test_vacuum_compacts_the_device_store_and_keeps_row_ids(gpu, host, oracle_mod):
    """LanceIndex::Vacuum (src/lance_index.cpp:728-751): after deletes, the rows are dropped from HBM for good and
    searches, filtered searches and later appends still speak the table's row ids."""
    rng = np.random.default_rng(51)
    n, d, k = 4000, 16, 8
    X = rng.standard_normal((n, d)).astype(np.float32)
    row_ids = np.arange(n, dtype=np.int64) * 3 + 7
    score = rng.integers(0, 100, n)
    with host.HostIndex(d, "l2") as ix:
        ix.append(X, row_ids)
        ix.columns.append("score", [int(v) for v in score])
        dead = np.sort(rng.choice(n, 1500, replace=False))
        ix.delete(row_ids[dead])
        keep = np.setdiff1d(np.arange(n), dead)
        q = rng.standard_normal(d).astype(np.float32)
        before, bd = ix.search(q, k)
        fb, _ = ix.search(q, k, "score < 40")
        ix.vacuum()
        assert ix.count() == len(keep)
        after, ad = ix.search(q, k)
        assert np.array_equal(after, before) and np.allclose(ad, bd, rtol=1e-6)
        wl, _, _ = oracle_mod.flat_search(X[keep], q, k, "l2")
        assert np.array_equal(after, row_ids[keep[wl]])
        fa, _ = ix.search(q, k, "score < 40")
        assert np.array_equal(fa, fb) and (score[(fa - 7) // 3] < 40).all()
        more = rng.standard_normal((5, d)).astype(np.float32)
        ix.append(more, np.arange(5, dtype=np.int64) + 10**6)
        ix.columns.append("score", [1] * 5)
        got, _ = ix.search(more[2], 1)
        assert got[0] == 10**6 + 2
        ix.vacuum()  # nothing pending: a no-op

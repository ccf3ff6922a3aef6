"""Pins oracle/ against every golden vector the reference holds for the kNN path:
the toy fixtures of test/sql/*.test, DuckDB's array_* known answers, and - where oracle/_ref was
built from the reference's own header - the reference's distance kernels on random vectors."""
import json
import math
import os

import numpy as np
import pytest

from parity import assert_topk_parity

HERE = os.path.dirname(os.path.abspath(__file__))
FX = json.load(open(os.path.join(HERE, "golden", "reference_sql_fixtures.json")))

# the predicate strings of lance_optimizer_filter.test, evaluated over the fixture's extra columns
PREDICATES = {
    "lang = 'en'": lambda c: np.array([v == "en" for v in c["lang"]]),
    "lang = 'es'": lambda c: np.array([v == "es" for v in c["lang"]]),
    "score > 20": lambda c: np.array([v > 20 for v in c["score"]]),
    "lang IS NOT NULL": lambda c: np.array([v is not None for v in c["lang"]]),
    "lang IN ('en', 'fr')": lambda c: np.array([v in ("en", "fr") for v in c["lang"]]),
    "NOT (lang = 'en')": lambda c: np.array([v != "en" for v in c["lang"]]),
}


def run_case(oracle_mod, c):
    X = np.array(c["rows"], np.float32)
    ids = np.array(c["ids"])
    allow = deleted = None
    if "predicate" in c:
        allow = oracle_mod.make_bitmap(PREDICATES[c["predicate"]](c["columns"]))
    if "deleted_ids" in c:
        deleted = oracle_mod.make_bitmap(np.isin(ids, c["deleted_ids"]))
    labels, d32, _ = oracle_mod.flat_search(X, np.array(c["query"], np.float32), c["k"], c["metric"], allow, deleted)
    return ids, labels, d32


@pytest.mark.parametrize("case", FX["cases"], ids=[c["name"] for c in FX["cases"]])
def test_reference_sql_fixture(oracle_mod, case):
    ids, labels, d32 = run_case(oracle_mod, case)
    if "expect_ids" in case:
        # tie-aware: lance_optimizer_filter.test:35-44 pins DuckDB's arbitrary pick between two rows at
        # exactly 2.0; BASELINE.md section 5 exempts ties, everything else must match id for id
        X = np.array(case["rows"], np.float64)
        q = np.array(case["query"], np.float64)
        truth = lambda lab: float(((X[lab] - q) ** 2).sum())
        want_labels = [list(ids).index(i) for i in case["expect_ids"]]
        assert_topk_parity(labels, d32, want_labels, [truth(l) for l in want_labels], truth, what=case["name"])
    if "expect_row_ids" in case:
        assert list(labels) == case["expect_row_ids"]
    if "expect_count" in case:
        assert len(labels) == case["expect_count"]
    if "expect_distances" in case:
        assert ["%.6f" % v for v in d32] == ["%.6f" % v for v in case["expect_distances"]]
    if "expect_ids_sorted_by_distance_then_any" in case:
        pos = 0
        for group in case["expect_ids_sorted_by_distance_then_any"]:
            assert sorted(ids[labels[pos:pos + len(group)]]) == sorted(group)
            pos += len(group)


def sql_float(v):
    """sqllogictest prints FLOAT with up to 7-8 significant digits; compare at that precision."""
    return float(np.float32(v))


def test_duckdb_known_answers(oracle_mod):
    ka = FX["duckdb_known_answers"]
    q = np.array(ka["array_distance"]["query"], np.float32)
    for row, want in zip(ka["array_distance"]["rows"], ka["array_distance"]["expect"]):
        got = math.sqrt(oracle_mod.distance_f32("l2", q, np.array(row, np.float32)))  # L2^2 = array_distance^2
        assert abs(got - want) <= 5e-7 * max(1.0, abs(want))
    for row, want in zip(ka["array_cosine_distance"]["rows"], ka["array_cosine_distance"]["expect"]):
        got = oracle_mod.distance_f32("cosine", q, np.array(row, np.float32))
        assert abs(got - want) <= 1e-7 + 2e-6 * abs(want)
    for row, want in zip(ka["array_inner_product"]["rows"], ka["array_inner_product"]["expect"]):
        got = 1.0 - oracle_mod.distance_f32("dot", q, np.array(row, np.float32))  # Lance dot = 1 - ip
        assert got == want


def test_against_reference_kernels(oracle_mod):
    """oracle/_ref = the reference's DuckDB kernels compiled from /root/reference (array_kernels.hpp)."""
    ref = oracle_mod.ref()
    if ref is None:
        pytest.skip("oracle/_ref not built (reference tree absent)")
    rng = np.random.default_rng(7)
    for d in (1, 3, 8, 128, 384, 768, 1000):
        for _ in range(20):
            a = rng.standard_normal(d).astype(np.float32)
            b = rng.standard_normal(d).astype(np.float32)
            # bit-exact: same sequential fp32 fold
            assert oracle_mod.distance_f32("l2", a, b) == ref.ref_array_distance_squared(a, b, d)
            assert np.float32(1.0) - np.float32(ref.ref_array_inner_product(a, b, d)) == np.float32(
                oracle_mod.distance_f32("dot", a, b))
            got, want = oracle_mod.distance_f32("cosine", a, b), ref.ref_array_cosine_distance(a, b, d)
            assert abs(got - want) <= 4e-7  # sqrt(nl*nr) vs one sqrt: last-ulp only
            # the fp64 truth agrees with the fp32 fold to fp32 accuracy
            for m in ("l2", "cosine", "dot"):
                t, f = oracle_mod.distance_f64(m, a, b), oracle_mod.distance_f32(m, a, b)
                assert abs(t - f) <= 1e-4 * max(abs(t), 1.0)


def test_flat_truth_vs_numpy(oracle_mod):
    rng = np.random.default_rng(11)
    X = rng.standard_normal((5000, 48)).astype(np.float32)
    Q = rng.standard_normal((4, 48)).astype(np.float32)
    Xd, Qd = X.astype(np.float64), Q.astype(np.float64)
    for metric in ("l2", "cosine", "dot"):
        for qi in range(4):
            if metric == "l2":
                dist = ((Xd - Qd[qi]) ** 2).sum(1)
            elif metric == "dot":
                dist = 1.0 - Xd @ Qd[qi]
            else:
                dist = 1.0 - (Xd @ Qd[qi]) / (np.linalg.norm(Xd, axis=1) * np.linalg.norm(Qd[qi]))
            order = np.lexsort((np.arange(len(dist)), dist))[:25]
            labels, d32, d64 = oracle_mod.flat_search(X, Q[qi], 25, metric)
            assert list(labels) == list(order)
            np.testing.assert_allclose(d64, dist[order], rtol=1e-12, atol=1e-12)


def test_ties_break_by_label_and_masks(oracle_mod):
    X = np.zeros((70, 4), np.float32)
    X[:, 0] = 1.0  # 70 exact duplicates
    q = np.array([1, 0, 0, 0], np.float32)
    labels, d32, _ = oracle_mod.flat_search(X, q, 5, "l2")
    assert list(labels) == [0, 1, 2, 3, 4] and not d32.any()
    deleted = oracle_mod.make_bitmap(np.arange(70) < 3)
    allow = oracle_mod.make_bitmap(np.arange(70) % 2 == 0)
    labels, _, _ = oracle_mod.flat_search(X, q, 4, "l2", allow, deleted)
    assert list(labels) == [4, 6, 8, 10]
    labels, _, _ = oracle_mod.flat_search(X, q, 100, "l2", allow, deleted)
    assert len(labels) == 33  # fewer than k survive
    labels, _, _ = oracle_mod.flat_search(X[:0], q, 3, "l2")
    assert len(labels) == 0  # empty table


def test_batch_f32_matches_truth(oracle_mod):
    rng = np.random.default_rng(3)
    X = rng.standard_normal((20000, 64)).astype(np.float32)
    Q = rng.standard_normal((9, 64)).astype(np.float32)
    deleted = oracle_mod.make_bitmap(rng.random(20000) < 0.1)
    for metric in ("l2", "cosine", "dot"):
        L, D, cnt = oracle_mod.flat_search_batch_f32(X, Q, 10, metric, None, deleted, nthreads=4)
        for qi in range(9):
            labels, d32, d64 = oracle_mod.flat_search(X, Q[qi], 10, metric, None, deleted)
            assert cnt[qi] == 10
            assert list(L[qi]) == list(labels)  # N(0,1) data: no near-ties at fp32 resolution here
            np.testing.assert_allclose(D[qi], d32, rtol=1e-5, atol=1e-6)


def test_merge_topk(oracle_mod):
    labels = np.array([[0, 5, 9], [3, 4, -1]], np.int64)
    dists = np.array([[0.1, 0.5, 0.9], [0.1, 0.6, 0.0]], np.float32)
    ol, od = oracle_mod.merge_topk(labels, dists, np.array([3, 2], np.int32), 3)
    assert list(ol) == [0, 3, 5]
    np.testing.assert_allclose(od, [0.1, 0.1, 0.5])


def test_ivfpq_oracle_self_consistency(oracle_mod):
    """No reference vectors exist for IVF_PQ (parity unpinned, oracle.h): check the restatement's
    internal contracts - probe order, LUT vs decoded residual, ADC ranking, refine = exact."""
    rng = np.random.default_rng(5)
    n, d, P, M = 3000, 32, 16, 8
    X = rng.standard_normal((n, d)).astype(np.float32)
    q = rng.standard_normal(d).astype(np.float32)
    for metric in ("l2", "cosine", "dot"):
        ix = oracle_mod.IvfPqIndex.train(X, metric, P, M, iters=5)
        assert ix.part_offsets[-1] == n and sorted(ix.row_ids) == list(range(n))
        parts, pd = ix.probe(q, 4)
        assert list(pd) == sorted(pd) and len(set(parts)) == 4
        # LUT entry == distance between the query residual and the codeword
        lut = ix.lut(q, parts[0])
        qq = q / np.linalg.norm(q) if metric == "cosine" else q
        r = qq - ix.centroids[parts[0]] if metric != "dot" else qq
        m, c = 3, 77
        sub = r[m * 4:(m + 1) * 4].astype(np.float64)
        cw = ix.codebook[m, c].astype(np.float64)
        want = -(sub @ cw) if metric == "dot" else ((sub - cw) ** 2).sum()
        assert abs(lut[m, c] - want) <= 1e-6 * max(1.0, abs(want))
        # all partitions probed + big refine == exact flat search
        res = ix.search(X, q, 10, nprobes=P, refine_factor=n // 10)
        labels, d32, _ = oracle_mod.flat_search(X, q, 10, metric)
        assert list(res["labels"]) == list(labels)
        np.testing.assert_allclose(res["dist"], d32, rtol=1e-6, atol=1e-6)
        # ADC candidates are sorted and come from the probed partitions only
        res = ix.search(X, q, 10, nprobes=4, refine_factor=3)
        assert len(res["cand_labels"]) == 30 and list(res["cand_adc"]) == sorted(res["cand_adc"])
        probed_rows = set()
        for p in res["parts"]:
            probed_rows |= set(ix.row_ids[ix.part_offsets[p]:ix.part_offsets[p + 1]])
        assert set(res["cand_labels"]) <= probed_rows
        # ADC approximates the exact distance
        exact = np.array([oracle_mod.distance_f64(metric, q, X[i]) for i in res["cand_labels"]])
        assert np.corrcoef(exact, res["cand_adc"])[0, 1] > 0.5

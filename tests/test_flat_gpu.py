"""GPU parity tests of the flat kNN path, through the C ABI (include/lance_b200.h), against the oracle.

Bar: returned ids identical to the oracle's except where truth distances tie within
1e-5 relative; distances within 1e-4 relative.  Squared L2 (test/sql/lance_basic.test:40-42)."""
import json
import os

import numpy as np
import pytest

from parity import assert_topk_parity

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
FX = json.load(open(os.path.join(HERE, "golden", "reference_sql_fixtures.json")))

PREDICATES = {
    "lang = 'en'": lambda c: np.array([v == "en" for v in c["lang"]]),
    "lang = 'es'": lambda c: np.array([v == "es" for v in c["lang"]]),
    "score > 20": lambda c: np.array([v > 20 for v in c["score"]]),
    "lang IS NOT NULL": lambda c: np.array([v is not None for v in c["lang"]]),
    "lang IN ('en', 'fr')": lambda c: np.array([v in ("en", "fr") for v in c["lang"]]),
    "NOT (lang = 'en')": lambda c: np.array([v != "en" for v in c["lang"]]),
}


def truth_fn(oracle_mod, X, q, metric):
    return lambda lab: oracle_mod.distance_f64(metric, q, X[lab])


def check_against_oracle(oracle_mod, h, X, Q, k, metric, allow=None, deleted=None, paths=(1,), what=""):
    """Single-query calls and one batch call, both compared with the oracle's truth ranking."""
    Q = np.atleast_2d(Q)
    for path in paths:
        h.set_flat_path(path)
        L, D, C = h.search_batch(Q, k, allow=allow)
        for j in range(Q.shape[0]):
            wl, wd32, wd64 = oracle_mod.flat_search(X, Q[j], k, metric, allow, deleted)
            assert C[j] == len(wl), f"{what} q{j}: count {C[j]} != {len(wl)}"
            assert_topk_parity(L[j, :C[j]], D[j, :C[j]], wl, wd64, truth_fn(oracle_mod, X, Q[j], metric),
                               what=f"{what} path{path} q{j}")
            assert (L[j, C[j]:] == -1).all() and np.isinf(D[j, C[j]:]).all()
            if j < 2:
                l1, d1 = h.search(Q[j], k, allow=allow)
                assert list(l1) == list(L[j, :C[j]])
                np.testing.assert_array_equal(d1, D[j, :C[j]])
    h.set_flat_path(0)


# ------------------------------------------------------------------ the reference's own golden vectors
@pytest.mark.parametrize("case", FX["cases"], ids=[c["name"] for c in FX["cases"]])
def test_reference_sql_fixture_on_gpu(gpu, oracle_mod, case):
    X = np.array(case["rows"], np.float32)
    ids = np.array(case["ids"])
    allow = None
    with gpu.LanceHandle(X.shape[1], case["metric"]) as h:
        labels = h.add_batch(X)
        assert list(labels) == list(range(len(X)))  # consecutive labels (lance_manager.rs:232)
        if "predicate" in case:
            allow = gpu.make_bitmap(PREDICATES[case["predicate"]](case["columns"]))
        if "deleted_ids" in case:
            h.delete_batch(np.nonzero(np.isin(ids, case["deleted_ids"]))[0])
            assert h.count() == len(X) - len(case["deleted_ids"])
        got_l, got_d = h.search(np.array(case["query"], np.float32), case["k"], allow=allow)
    if "expect_ids" in case:
        Xd, q = X.astype(np.float64), np.array(case["query"], np.float64)
        truth = lambda lab: float(((Xd[lab] - q) ** 2).sum())
        want = [list(ids).index(i) for i in case["expect_ids"]]
        assert_topk_parity(got_l, got_d, want, [truth(l) for l in want], truth, what=case["name"])
    if "expect_row_ids" in case:
        assert list(got_l) == case["expect_row_ids"]
    if "expect_count" in case:
        assert len(got_l) == case["expect_count"]
    if "expect_distances" in case:
        assert ["%.6f" % v for v in got_d] == ["%.6f" % v for v in case["expect_distances"]]
    if "expect_ids_sorted_by_distance_then_any" in case:
        pos = 0
        for group in case["expect_ids_sorted_by_distance_then_any"]:
            assert sorted(ids[got_l[pos:pos + len(group)]]) == sorted(group)
            pos += len(group)


def test_duckdb_known_answers_on_gpu(gpu):
    """DuckDB's array_* known answers (duckdb/test/sql/function/array/*.test), via Lance conventions."""
    ka = FX["duckdb_known_answers"]
    q = np.array(ka["array_distance"]["query"], np.float32)
    for name, metric, conv in (("array_distance", "l2", lambda d: np.sqrt(d)),
                               ("array_cosine_distance", "cosine", lambda d: d),
                               ("array_inner_product", "dot", lambda d: 1.0 - d)):
        rows = np.array(ka[name]["rows"], np.float32)
        with gpu.LanceHandle(rows.shape[1], metric) as h:
            h.add_batch(rows)
            labels, dist = h.search(q, len(rows))
        got = np.empty(len(rows))
        got[labels] = conv(dist.astype(np.float64))
        np.testing.assert_allclose(got, ka[name]["expect"], rtol=2e-6, atol=2e-7)


# ------------------------------------------------------------------ seeded random parity
@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
@pytest.mark.parametrize("n,d,k", [(100_000, 128, 10), (20_000, 768, 100), (5_003, 3, 7), (4_099, 30, 33),
                                   (33_333, 384, 10)])
def test_flat_random_parity(gpu, oracle_mod, metric, n, d, k):
    rng = np.random.default_rng(0xC0FFEE + n + d)
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((5, d), dtype=np.float32)
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X[: n // 2])
        h.add_batch(X[n // 2:])  # growth path
        assert h.count() == n
        check_against_oracle(oracle_mod, h, X, Q, k, metric, what=f"{metric} {n}x{d} k={k}")
        st = h.stats()
        assert st["kernel_launches"] > 0 and st["flat_simt_batches"] > 0


@pytest.mark.parametrize("nq", [1, 2, 3, 4, 5, 9])
def test_query_batch_sizes(gpu, oracle_mod, nq):
    rng = np.random.default_rng(nq)
    X = rng.standard_normal((30_000, 64), dtype=np.float32)
    Q = rng.standard_normal((nq, 64), dtype=np.float32)
    with gpu.LanceHandle(64, "l2") as h:
        h.add_batch(X)
        check_against_oracle(oracle_mod, h, X, Q, 10, "l2", what=f"nq={nq}")


@pytest.mark.parametrize("d,nq,k", [(3, 1, 3), (5, 3, 7), (7, 5, 1), (33, 2, 11), (129, 9, 5)])
def test_host_buffer_results_with_odd_dimensions(gpu, oracle_mod, d, nq, k):
    """Small host-buffer calls get their packed [labels | distances | counts] result written in place into the
    mapped staging buffer behind the queries (store.cu search_host): the int64 labels must stay aligned whatever
    nq * d * 4 is, and the caller's buffers must hold exactly what the device path returns."""
    rng = np.random.default_rng(100 * d + nq)
    X = rng.standard_normal((4000, d), dtype=np.float32)
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    for metric in ("l2", "cosine", "dot"):
        with gpu.LanceHandle(d, metric) as h:
            h.add_batch(X)
            check_against_oracle(oracle_mod, h, X, Q, k, metric, what=f"d={d} nq={nq} k={k} {metric}")
            # repeated calls reuse the staging buffer: same answer every time
            a = h.search(Q[0], k)
            b = h.search(Q[0], k)
            assert list(a[0]) == list(b[0]) and np.array_equal(a[1], b[1])


def test_edge_cases(gpu, oracle_mod):
    rng = np.random.default_rng(5)
    d = 16
    X = rng.standard_normal((1000, d), dtype=np.float32)
    q = rng.standard_normal(d).astype(np.float32)
    with gpu.LanceHandle(d, "l2") as h:
        # empty table: zero results, no error (lance_delete.test k>n semantics extended to n=0)
        l, dist = h.search(q, 5)
        assert len(l) == 0
        h.add_batch(X[:3])
        l, dist = h.search(q, 10)  # k > n returns n (lance_delete.test:22-34)
        wl, _, wd = oracle_mod.flat_search(X[:3], q, 10, "l2")
        assert list(l) == list(wl)
        l, dist = h.search(q, 0)
        assert len(l) == 0
        h.add_batch(X[3:])
        for k in (1, 2, 31, 32, 33, 64, 100, 257, 1000, 1500):
            check_against_oracle(oracle_mod, h, X, q, k, "l2", what=f"k={k}")
        # wrong dimension is an error at this level (lance_manager.rs:401-407) ...
        with pytest.raises(gpu.LanceError, match="expected query dimension 16, got 8"):
            h.search(q[:8], 3)
        with pytest.raises(gpu.LanceError, match="expected dimension"):
            h.add_batch(np.zeros((2, 5), np.float32))
        # ... and an HNSW request (rust_ffi.cpp:35-36) is accepted without building a graph: searches stay exact
        h.create_hnsw_index(20, 50)
        h.create_hnsw_index(0, 0)
        assert h.stats()["has_index"] == 0
        check_against_oracle(oracle_mod, h, X, q, 5, "l2", what="after create_hnsw_index")
        assert h.count() == 1000
        np.testing.assert_array_equal(h.get_vector(17), X[17])


def test_ties_break_by_label_and_masks(gpu, oracle_mod):
    """Exact duplicates: ascending label decides (the oracle's tie rule); prefilter + tombstones compose."""
    X = np.zeros((70, 4), np.float32)
    X[:, 0] = 1.0
    q = np.array([1, 0, 0, 0], np.float32)
    with gpu.LanceHandle(4, "l2") as h:
        h.add_batch(X)
        l, d = h.search(q, 5)
        assert list(l) == [0, 1, 2, 3, 4] and not d.any()
        h.delete_batch([0, 1, 2])
        h.delete_batch([2, 500, -3])  # repeated / unknown labels match nothing
        assert h.count() == 67
        allow = gpu.make_bitmap(np.arange(70) % 2 == 0)
        l, _ = h.search(q, 4, allow=allow)
        assert list(l) == [4, 6, 8, 10]
        l, _ = h.search(q, 100, allow=allow)
        assert len(l) == 33  # fewer than k survive (lance_optimizer_filter.test:56-64)
        # a bitmap shorter than the table filters the uncovered labels out
        l, _ = h.search(q, 100, allow=gpu.make_bitmap(np.ones(32, bool)))
        assert list(l) == list(range(3, 32))
        labels, vecs = h.get_all_vectors()
        assert list(labels) == list(range(3, 70)) and (vecs == X[3:]).all()


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_prefilter_and_deletes_random(gpu, oracle_mod, metric):
    rng = np.random.default_rng(21)
    n, d = 50_000, 96
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((4, d), dtype=np.float32)
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X)
        dead = rng.choice(n, 5000, replace=False)
        h.delete_batch(dead)
        deleted = oracle_mod.make_bitmap(np.isin(np.arange(n), dead))
        for sel in (0.8, 0.01, 0.0002):  # BASELINE config 5's two selectivities and a near-empty one
            allow = gpu.make_bitmap(rng.random(n) < sel)
            check_against_oracle(oracle_mod, h, X, Q, 10, metric, allow, deleted, what=f"{metric} sel={sel}")
        check_against_oracle(oracle_mod, h, X, Q, 10, metric, None, deleted, what=f"{metric} deletes only")


def test_zero_vectors_and_nan_free_cosine(gpu, oracle_mod):
    """cosine with a zero row / zero query is distance 1 (array_kernels.hpp clamp convention in the oracle)."""
    rng = np.random.default_rng(2)
    X = rng.standard_normal((500, 8), dtype=np.float32)
    X[7] = 0
    with gpu.LanceHandle(8, "cosine") as h:
        h.add_batch(X)
        q = rng.standard_normal(8).astype(np.float32)
        check_against_oracle(oracle_mod, h, X, q, 500, "cosine", what="zero row")
        l, d = h.search(np.zeros(8, np.float32), 3)
        assert list(l) == [0, 1, 2] and np.allclose(d, 1.0)


def test_label_base_and_device_entry(gpu, oracle_mod):
    """Shard semantics: labels start at the shard's base; device-resident search + device merge of two shards
    equal the unsharded search (top-k of a union = top-k of per-shard top-k's)."""
    import torch

    rng = np.random.default_rng(9)
    n, d, k, nq = 40_000, 128, 10, 3
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    dev = torch.device("cuda:0")
    dQ = torch.from_numpy(Q).to(dev)
    cut = 17_001
    outs = []
    handles = []
    for base, part in ((0, X[:cut]), (cut, X[cut:])):
        h = gpu.LanceHandle(d, "l2")
        h.set_label_base(base)
        dX = torch.from_numpy(part).to(dev)
        first = h.add_batch_device(dX.data_ptr(), part.shape[0], d)
        assert first == base
        L = torch.empty((nq, k), dtype=torch.int64, device=dev)
        D = torch.empty((nq, k), dtype=torch.float32, device=dev)
        C = torch.empty(nq, dtype=torch.int32, device=dev)
        h.search_device(dQ.data_ptr(), nq, k, L.data_ptr(), D.data_ptr(), C.data_ptr(),
                        stream=torch.cuda.current_stream().cuda_stream)
        outs.append((L, D, C))
        handles.append(h)
    gl = torch.stack([o[0] for o in outs]).contiguous()
    gd = torch.stack([o[1] for o in outs]).contiguous()
    gc = torch.stack([o[2] for o in outs]).contiguous()
    ML = torch.empty((nq, k), dtype=torch.int64, device=dev)
    MD = torch.empty((nq, k), dtype=torch.float32, device=dev)
    MC = torch.empty(nq, dtype=torch.int32, device=dev)
    gpu.merge_topk_device(0, gl.data_ptr(), gd.data_ptr(), gc.data_ptr(), 2, nq, k, ML.data_ptr(), MD.data_ptr(),
                          MC.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for j in range(nq):
        wl, _, wd = oracle_mod.flat_search(X, Q[j], k, "l2")
        assert MC[j].item() == k
        assert_topk_parity(ML[j].cpu().numpy(), MD[j].cpu().numpy(), wl, wd, truth_fn(oracle_mod, X, Q[j], "l2"),
                           what=f"merged q{j}")
        # and against the checker's own merge of the two shard lists
        ol, od = oracle_mod.merge_topk(gl[:, j].cpu().numpy(), gd[:, j].cpu().numpy(), gc[:, j].cpu().numpy(), k)
        assert list(ol) == list(ML[j].cpu().numpy())
    for h in handles:
        h.close()


def test_save_and_open(gpu, oracle_mod, tmp_path):
    """lance_open_detached (rust_lib/src/ffi.rs:93-111) re-opens what the store saved; deletes persist
    (test/sql/lance_persist.test:20-34, lance_delete.test:37-46)."""
    rng = np.random.default_rng(4)
    X = rng.standard_normal((3000, 24), dtype=np.float32)
    q = rng.standard_normal(24).astype(np.float32)
    with gpu.LanceHandle(24, "l2", db_path=str(tmp_path), table_name="t_idx") as h:
        h.add_batch(X)
        h.delete_batch([5, 6])
        want = h.search(q, 10)
        h.save()
    with gpu.LanceHandle.open(str(tmp_path), "t_idx", "l2") as h2:
        assert h2.dim == 24 and h2.count() == 2998
        got = h2.search(q, 10)
        assert list(got[0]) == list(want[0])
        new = h2.add_batch(X[:2])  # labels continue after restart (lance_manager.rs:157-158)
        assert list(new) == [3000, 3001]
    with pytest.raises(gpu.LanceError, match="open failed"):
        gpu.LanceHandle.open(str(tmp_path), "missing", "l2")


def test_concurrent_searches_one_handle(gpu, oracle_mod):
    """The C ABI is re-entrant on one handle (SURVEY.md 8b threading): 4 threads search while one appends."""
    import threading

    rng = np.random.default_rng(8)
    X = rng.standard_normal((20_000, 32), dtype=np.float32)
    Q = rng.standard_normal((16, 32), dtype=np.float32)
    extra = rng.standard_normal((4096, 32), dtype=np.float32) + 100.0  # far away: never in any top-k
    with gpu.LanceHandle(32, "l2") as h:
        h.add_batch(X)
        want = [oracle_mod.flat_search(X, Q[j], 10, "l2")[0] for j in range(16)]
        errors = []

        def worker(t):
            try:
                for rep in range(20):
                    j = (t * 5 + rep) % 16
                    l, _ = h.search(Q[j], 10)
                    assert list(l) == list(want[j]), (t, rep)
            except Exception as e:  # noqa: BLE001
                errors.append(e)

        def writer():
            try:
                for i in range(0, 4096, 512):
                    h.add_batch(extra[i:i + 512])
            except Exception as e:  # noqa: BLE001
                errors.append(e)

        ts = [threading.Thread(target=worker, args=(t,)) for t in range(4)] + [threading.Thread(target=writer)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        assert not errors, errors
        assert h.count() == 20_000 + 4096


def test_merge_moves_live_rows_between_handles(gpu, oracle_mod):
    """lance_detached_merge (rust_lib/src/ffi.rs:187-228, lance_manager.rs:307-393): the listed live rows of the source
    land in the target under fresh sequential labels, in the source's storage order; unknown and deleted labels are
    skipped; searching the target equals searching the concatenated table."""
    rng = np.random.default_rng(808)
    d = 24
    A = rng.standard_normal((700, d), dtype=np.float32)
    B = rng.standard_normal((500, d), dtype=np.float32)
    with gpu.LanceHandle(d, "l2") as ta, gpu.LanceHandle(d, "l2") as sb:
        ta.add_batch(A)
        sb.add_batch(B)
        sb.delete_batch([3, 77, 499])
        want = [498, 5, 3, 10_000, -4, 77, 250, 5, 0]  # unsorted, a duplicate, deleted and unknown labels
        old, new = ta.merge(sb, want)
        assert old.tolist() == [0, 5, 250, 498]
        assert new.tolist() == [700, 701, 702, 703]
        assert ta.count() == 704
        for o, n in zip(old, new):
            np.testing.assert_array_equal(ta.get_vector(int(n)), B[o])
        X = np.concatenate([A, B[old]])
        for q in (B[250], A[13], rng.standard_normal(d).astype(np.float32)):
            l, dd = ta.search(q, 7)
            wl, wd32, wd64 = oracle_mod.flat_search(X, q, 7, "l2")
            assert list(l) == list(wl)
            np.testing.assert_allclose(dd, wd64, rtol=1e-4, atol=1e-6)
        # nothing to merge, and the error contract
        o2, n2 = ta.merge(sb, [])
        assert len(o2) == 0 and ta.count() == 704
        with pytest.raises(gpu.LanceError, match="merge"):
            ta.merge(ta, [1])
        with gpu.LanceHandle(d + 1, "l2") as other:
            other.add_batch(rng.standard_normal((4, d + 1), dtype=np.float32))
            with pytest.raises(gpu.LanceError, match="dimension mismatch"):
                ta.merge(other, [0])

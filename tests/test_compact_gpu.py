"""lance_detached_compact drops tombstoned rows for good and keeps labels (f1: src/lance_index.cpp:728-751 Vacuum ->
LanceDetachedCompact -> lance_manager.rs:553-566), explicit labels (lance_b200_add_batch_labeled), through the C ABI."""
import numpy as np
import pytest

from parity import assert_topk_parity

pytestmark = pytest.mark.gpu


def _check(oracle_mod, h, X, labels_of_rows, q, k, metric, mask=None, allow=None, what=""):
    """h's result vs the oracle over the rows X (whose labels are labels_of_rows)."""
    a = None if mask is None else oracle_mod.make_bitmap(mask)
    wl, _, wd = oracle_mod.flat_search(X, q, k, metric, a)
    gl, gd = h.search(q, k, allow=allow)
    row_of = {int(l): i for i, l in enumerate(labels_of_rows)}
    assert_topk_parity(gl, gd, labels_of_rows[wl], wd, lambda lab: oracle_mod.distance_f64(metric, q, X[row_of[lab]]), what=what)


@pytest.mark.parametrize("metric,n", [("l2", 5000), ("cosine", 70000)])
def test_compact_keeps_labels_and_results(gpu, oracle_mod, metric, n):
    rng = np.random.default_rng(41)
    d, k = 32, 10
    X = rng.standard_normal((n, d)).astype(np.float32)
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X)
        dead = np.sort(rng.choice(n, n // 3, replace=False))
        h.delete_batch(dead.astype(np.int64))
        keep = np.setdiff1d(np.arange(n), dead)
        Q = rng.standard_normal((12, d)).astype(np.float32)
        before = [h.search(q, k) for q in Q[:3]]
        h.compact()
        st = h.stats()
        assert st["rows"] == len(keep) and st["live_rows"] == len(keep)  # the rows are gone, not just hidden
        for q, (bl, bd) in zip(Q[:3], before):
            al, ad = h.search(q, k)
            assert np.array_equal(al, bl) and np.allclose(ad, bd, rtol=1e-6)  # same labels as before the compaction
            _check(oracle_mod, h, X[keep], keep, q, k, metric, what="after compact")
        L, D, C = h.search_batch(Q, k)  # batched path (tcgen05 when the table is large enough)
        for i, q in enumerate(Q):
            wl, _, wd = oracle_mod.flat_search(X[keep], q, k, metric)
            assert_topk_parity(L[i, :C[i]], D[i, :C[i]], keep[wl], wd,
                               lambda lab, q=q: oracle_mod.distance_f64(metric, q, X[lab]), what=f"batch q{i}")
        # labels keep working at the boundary: get_vector, delete, label-indexed allow bitmap, append
        lab = int(keep[7])
        assert np.array_equal(h.get_vector(lab), X[lab])
        with pytest.raises(gpu.LanceError, match="not found"):
            h.get_vector(int(dead[0]))
        h.delete_batch(np.array([lab, int(dead[0])], np.int64))  # a dropped label matches nothing
        assert h.count() == len(keep) - 1
        keep2 = keep[keep != lab]
        mask_by_label = np.zeros(n, bool)
        mask_by_label[keep2[::2]] = True
        allow = gpu.make_bitmap(mask_by_label)
        _check(oracle_mod, h, X[keep2], keep2, Q[0], k, metric, mask=mask_by_label[keep2], allow=allow, what="allow by label")
        new = rng.standard_normal((100, d)).astype(np.float32)
        nl = h.add_batch(new)
        assert np.array_equal(nl, np.arange(n, n + 100))  # the counter never reuses a label
        allX = np.concatenate([X[keep2], new])
        alll = np.concatenate([keep2, nl])
        _check(oracle_mod, h, allX, alll, Q[1], k, metric, what="after append")
        gl, gv = h.get_all_vectors()
        assert np.array_equal(gl, alll) and np.array_equal(gv, allX)


def test_compact_carries_columns_and_index(gpu, oracle_mod, tmp_path):
    rng = np.random.default_rng(42)
    n, d, k = 20000, 32, 10
    X = rng.standard_normal((n, d)).astype(np.float32)
    score = rng.integers(0, 100, n)
    db = str(tmp_path / "db")
    with gpu.LanceHandle(d, "l2", db, "t") as h:
        h.add_batch(X)
        h.set_column("score", score.astype(np.int64))
        h.set_column("lang", [["en", "fr", "es"][i % 3] for i in range(n)], type="str")
        h.create_index(16, 8)
        dead = np.sort(rng.choice(n, 5000, replace=False))
        h.delete_batch(dead.astype(np.int64))
        keep = np.setdiff1d(np.arange(n), dead)
        q = rng.standard_normal(d).astype(np.float32)
        bl, bd = h.search(q, k, nprobes=16)  # every partition probed
        pl, pd = h.search_predicate(q, k, "score < 30 AND lang = 'fr'", nprobes=16)
        h.compact()
        assert h.stats()["has_index"] == 1 and h.stats()["rows"] == len(keep)
        al, ad = h.search(q, k, nprobes=16)
        assert np.array_equal(al, bl) and np.allclose(ad, bd, rtol=1e-6)
        ql, qd = h.search_predicate(q, k, "score < 30 AND lang = 'fr'", nprobes=16)
        assert np.array_equal(ql, pl) and np.allclose(qd, pd, rtol=1e-6)
        assert (score[ql] < 30).all() and all(l % 3 == 1 for l in ql)
        h.save()
    with gpu.LanceHandle.open(db, "t", "l2") as h2:  # a compacted store survives a restart (labels, index, columns)
        assert h2.stats()["has_index"] == 1 and h2.count() == len(keep)
        rl, rd = h2.search(q, k, nprobes=16)
        assert np.array_equal(rl, bl) and np.allclose(rd, bd, rtol=1e-6)
        sl, sd = h2.search_predicate(q, k, "score < 30 AND lang = 'fr'", nprobes=16)
        assert np.array_equal(sl, pl)
        nl = h2.add_batch(rng.standard_normal((3, d)).astype(np.float32))
        assert np.array_equal(nl, [n, n + 1, n + 2])


def test_labeled_appends_with_gaps(gpu, oracle_mod):
    rng = np.random.default_rng(43)
    d, k = 16, 5
    chunks = [(1000, 300), (5000, 200), (5200, 100), (90000, 400)]  # (first label, rows): gaps between chunks
    with gpu.LanceHandle(d, "dot") as h:
        Xs, Ls = [], []
        for first, m in chunks:
            x = rng.standard_normal((m, d)).astype(np.float32)
            h.add_batch_labeled(x, first)
            Xs.append(x)
            Ls.append(np.arange(first, first + m))
        X, lab = np.concatenate(Xs), np.concatenate(Ls)
        assert h.count() == len(lab)
        for q in rng.standard_normal((3, d)).astype(np.float32):
            _check(oracle_mod, h, X, lab, q, k, "dot", what="gapped labels")
        with pytest.raises(gpu.LanceError, match="labels must ascend"):
            h.add_batch_labeled(X[:2], 5100)
        h.delete_batch(np.array([5000, 4000], np.int64))  # 4000 was never handed out
        assert h.count() == len(lab) - 1

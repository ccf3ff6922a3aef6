"""lance_b200_create_sharded: one handle over several devices inside ONE process, through the plain
lance_detached_* calls.  On a one-GPU box the shards share device 0 (the fan-out, label routing and merge are the same
code); with more GPUs visible every shard sits on its own device."""
import numpy as np
import pytest

from parity import assert_topk_parity

pytestmark = pytest.mark.gpu


def _devices(gpu, want):
    n = gpu.device_count()
    return [i % n for i in range(want)]


@pytest.mark.parametrize("metric,shards", [("l2", 2), ("dot", 3), ("cosine", 4)])
def test_sharded_handle_matches_oracle(gpu, oracle_mod, metric, shards):
    rng = np.random.default_rng(61)
    n, d, k = 30000, 48, 10
    X = rng.standard_normal((n, d)).astype(np.float32)
    h = gpu.LanceHandle.create_sharded(_devices(gpu, shards), d, metric)
    try:
        labels = []
        for lo in range(0, n, 2048):  # DuckDB hands the index 2048-row chunks (src/lance_index.cpp:941-946)
            labels.append(h.add_batch(X[lo:lo + 2048]))
        assert np.array_equal(np.concatenate(labels), np.arange(n))  # one global label sequence
        rows = h.shard_rows()
        assert len(rows) == shards and rows.sum() == n and rows.max() - rows.min() <= 2048
        assert h.count() == n and h.dim == d
        Q = rng.standard_normal((16, d)).astype(np.float32)
        truth = lambda q: (lambda lab: oracle_mod.distance_f64(metric, q, X[lab]))
        for q in Q[:4]:
            gl, gd = h.search(q, k)
            wl, _, wd = oracle_mod.flat_search(X, q, k, metric)
            assert_topk_parity(gl, gd, wl, wd, truth(q), what="single")
        L, D, C = h.search_batch(Q, k)
        for i, q in enumerate(Q):
            wl, _, wd = oracle_mod.flat_search(X, q, k, metric)
            assert_topk_parity(L[i, :C[i]], D[i, :C[i]], wl, wd, truth(q), what=f"batch q{i}")
        # deletes and a label-indexed prefilter are routed / replicated to the shards
        dead = rng.choice(n, 4000, replace=False)
        h.delete_batch(dead.astype(np.int64))
        mask = rng.random(n) < 0.3
        live = mask.copy()
        live[dead] = False
        gl, gd = h.search(Q[0], k, allow=gpu.make_bitmap(mask))
        wl, _, wd = oracle_mod.flat_search(X, Q[0], k, metric, oracle_mod.make_bitmap(live))
        assert_topk_parity(gl, gd, wl, wd, truth(Q[0]), what="filtered")
        assert np.array_equal(h.get_vector(int(np.flatnonzero(live)[5])), X[np.flatnonzero(live)[5]])
        h.compact()
        assert h.count() == n - 4000 and h.shard_rows().sum() == n - 4000
        alive = np.ones(n, bool)
        alive[dead] = False
        gl, gd = h.search(Q[1], k)
        wl, _, wd = oracle_mod.flat_search(X, Q[1], k, metric, oracle_mod.make_bitmap(alive))
        assert_topk_parity(gl, gd, wl, wd, truth(Q[1]), what="after compact")
    finally:
        h.close()


def test_sharded_handle_predicates_and_unsupported_calls(gpu, oracle_mod):
    rng = np.random.default_rng(62)
    n, d, k = 9000, 16, 6
    X = rng.standard_normal((n, d)).astype(np.float32)
    score = rng.integers(0, 100, n)
    lang = [["en", "fr", "es"][i % 3] for i in range(n)]
    h = gpu.LanceHandle.create_sharded(_devices(gpu, 3), d, "l2")
    try:
        for lo in range(0, n, 1000):
            h.add_batch(X[lo:lo + 1000])
        h.set_column("score", score.astype(np.int64))  # one call over the whole label range: cut at the batch boundaries
        h.set_column("lang", lang, type="str")
        assert h.has_extra_columns()
        q = rng.standard_normal(d).astype(np.float32)
        mask = (score < 25) & (np.arange(n) % 3 == 1)
        gl, gd = h.search_predicate(q, k, "score < 25 AND lang = 'fr'")
        wl, _, wd = oracle_mod.flat_search(X, q, k, "l2", oracle_mod.make_bitmap(mask))
        assert_topk_parity(gl, gd, wl, wd, lambda lab: oracle_mod.distance_f64("l2", q, X[lab]), what="predicate")
        with pytest.raises(gpu.LanceError, match="no db_path to save under"):
            h.save()  # an in-memory sharded handle (lance_b200_create_sharded); see test_sharded_handle_save_and_reopen
        with pytest.raises(gpu.LanceError, match="not supported on a sharded handle"):
            h.get_ivfpq()
        with pytest.raises(gpu.LanceError, match="expected query dimension"):
            h.search(np.zeros(d + 1, np.float32), k)
    finally:
        h.close()
    with pytest.raises(gpu.LanceError, match="does not exist"):
        gpu.LanceHandle.create_sharded([0, 99], d)


@pytest.mark.parametrize("metric", ["l2", "cosine"])
def test_sharded_handle_ivfpq_one_index_per_shard(gpu, oracle_mod, metric):
    """create_index on a sharded handle: every shard trains an index of ITS rows with the caller's parameters; a search
    probes nprobes lists on every shard and the shards' refined lists are merged.  (k-means sums with float atomics, so two
    trainings of the same rows are not bit-identical: the checks are properties of the merged answer, not equality with
    a second training.)"""
    rng = np.random.default_rng(63)
    n, d, k, P, M, nprobes = 40_000, 32, 10, 16, 8, 4
    centers = rng.standard_normal((64, d)).astype(np.float32) * 3
    X = (centers[rng.integers(0, 64, n)] + rng.standard_normal((n, d))).astype(np.float32)
    Q = (centers[rng.integers(0, 64, 24)] + rng.standard_normal((24, d))).astype(np.float32)
    half = n // 2
    h = gpu.LanceHandle.create_sharded(_devices(gpu, 2), d, metric)
    try:
        h.add_batch(X[:half])  # the least-filled shard takes a whole batch: shard 0, then shard 1
        h.add_batch(X[half:])
        assert list(h.shard_rows()) == [half, n - half]
        h.create_index(P, M)
        st0 = h.stats()
        assert st0["has_index"] == 1
        L, D, C = h.search_batch(Q, k, nprobes=nprobes, refine_factor=3)
        assert h.stats()["ivfpq_batches"] == st0["ivfpq_batches"] + 2, "both shards must answer from their index"
        hits = 0
        from_both = 0
        for i in range(Q.shape[0]):
            assert C[i] == k
            got = [int(x) for x in L[i]]
            assert len(set(got)) == k and all(0 <= g < n for g in got)
            # refined distances are exact distances of the returned rows, and the list is in (distance, label) order
            want_d = np.array([oracle_mod.distance_f64(metric, Q[i], X[g]) for g in got])
            np.testing.assert_allclose(D[i], want_d, rtol=1e-4, atol=1e-5)
            assert all((D[i, j], got[j]) <= (D[i, j + 1], got[j + 1]) for j in range(k - 1))
            from_both += int(min(got) < half <= max(got))
            wl, _, _ = oracle_mod.flat_search(X, Q[i], k, metric)
            hits += len(set(got) & set(int(x) for x in wl))
        assert from_both >= Q.shape[0] // 2, "results come from one shard only: the fan-out or the merge is broken"
        assert hits >= 0.5 * k * Q.shape[0], f"recall {hits / (k * Q.shape[0]):.2f} of the approximate search is implausibly low"
        # probing every list with a deep refine is the exact search
        Lx, Dx, Cx = h.search_batch(Q[:4], k, nprobes=P, refine_factor=50)
        for i in range(4):
            wl, _, wd = oracle_mod.flat_search(X, Q[i], k, metric)
            assert_topk_parity(Lx[i], Dx[i], wl, wd, lambda lab: oracle_mod.distance_f64(metric, Q[i], X[lab]), what=f"all lists q{i}")
        # an HNSW request replaces the index: exact again
        h.create_hnsw_index(16, 100)
        assert h.stats()["has_index"] == 0
        gl, gd = h.search(Q[0], k)
        wl, _, wd = oracle_mod.flat_search(X, Q[0], k, metric)
        assert_topk_parity(gl, gd, wl, wd, lambda lab: oracle_mod.distance_f64(metric, Q[0], X[lab]), what="after hnsw request")
    finally:
        h.close()


def test_sharded_handle_save_and_reopen(gpu, oracle_mod, tmp_path):
    """A sharded handle with a place on disk (lance_b200_create_sharded_at): every shard writes its store.bin, the manifest
    keeps the label counter and the batch -> shard map; the reopened handle answers like the one that was saved
    (test/sql/lance_persist.test semantics for an index spread over several GPUs)."""
    rng = np.random.default_rng(64)
    n, d, k = 12_000, 24, 8
    X = rng.standard_normal((n, d)).astype(np.float32)
    score = rng.integers(0, 50, n).astype(np.int64)
    Q = rng.standard_normal((6, d)).astype(np.float32)
    db = str(tmp_path / "db")
    h = gpu.LanceHandle.create_sharded(_devices(gpu, 3), d, "cosine", db_path=db, table="docs")
    dead = rng.choice(n, 900, replace=False).astype(np.int64)
    try:
        for lo in range(0, n, 1500):
            h.add_batch(X[lo:lo + 1500])
        h.set_column("score", score)
        h.delete_batch(dead)
        before = h.search_batch(Q, k)
        before_pred = h.search_batch_predicate(Q, k, "score >= 40")
        rows_before = h.shard_rows()
        h.save()
    finally:
        h.close()
    with pytest.raises(gpu.LanceError, match="saved with 3 shards"):
        gpu.LanceHandle.open_sharded(db, _devices(gpu, 2), "cosine", table="docs")
    with pytest.raises(gpu.LanceError, match="no saved sharded store"):
        gpu.LanceHandle.open_sharded(db, _devices(gpu, 3), "cosine", table="other")
    h = gpu.LanceHandle.open_sharded(db, _devices(gpu, 3), "cosine", table="docs")
    try:
        assert h.dim == d and h.count() == n - 900 and np.array_equal(h.shard_rows(), rows_before)
        after = h.search_batch(Q, k)
        after_pred = h.search_batch_predicate(Q, k, "score >= 40")
        for a, b in zip(before + before_pred, after + after_pred):
            np.testing.assert_array_equal(a, b)
        alive = np.ones(n, bool)
        alive[dead] = False
        wl, _, wd = oracle_mod.flat_search(X, Q[0], k, "cosine", oracle_mod.make_bitmap(alive))
        assert_topk_parity(after[0][0], after[1][0], wl, wd, lambda lab: oracle_mod.distance_f64("cosine", Q[0], X[lab]), what="reopened")
        # the label sequence goes on where it stopped, deletes still find their shard
        Y = rng.standard_normal((100, d)).astype(np.float32)
        first = h.add_batch(Y)
        assert int(np.atleast_1d(first)[0]) == n
        assert np.array_equal(h.get_vector(n + 7), Y[7])
        h.delete_batch(np.array([5 if 5 not in dead else int(np.flatnonzero(alive)[0])], np.int64))
        assert h.count() == n - 900 + 100 - 1
    finally:
        h.close()


def test_reference_entry_points_shard_by_environment(gpu, oracle_mod, tmp_path, monkeypatch):
    """LANCE_B200_DEVICES: the reference's own lance_create_detached / lance_open_detached hand out a sharded handle, so the
    unmodified extension spreads an index over the GPUs.  Includes MergeIndexes' calls (lance_detached_merge,
    lance_detached_get_all_vectors: src/lance_index.cpp:620-700) with a sharded side."""
    rng = np.random.default_rng(65)
    n, d, k = 6000, 20, 5
    X = rng.standard_normal((n, d)).astype(np.float32)
    q = rng.standard_normal(d).astype(np.float32)
    ndev = gpu.device_count()
    monkeypatch.setenv("LANCE_B200_DEVICES", ",".join(str(i % ndev) for i in range(3)))
    db = str(tmp_path / "db")
    h = gpu.LanceHandle(d, "l2", db_path=db, table_name="t")  # lance_create_detached
    try:
        assert len(h.shard_rows()) == 3
        for lo in range(0, n, 1000):
            h.add_batch(X[lo:lo + 1000])
        gl, gd = h.search(q, k)
        wl, _, wd = oracle_mod.flat_search(X, q, k, "l2")
        assert_topk_parity(gl, gd, wl, wd, lambda lab: oracle_mod.distance_f64("l2", q, X[lab]), what="env-sharded")
        labs, vecs = h.get_all_vectors()
        order = np.argsort(labs)
        assert np.array_equal(labs[order], np.arange(n)) and np.array_equal(vecs[order], X)
        h.save()
    finally:
        h.close()
    h = gpu.LanceHandle.open(db, "t", "l2")  # lance_open_detached finds the manifest
    monkeypatch.delenv("LANCE_B200_DEVICES")
    plain = gpu.LanceHandle(d, "l2")
    try:
        assert len(h.shard_rows()) == 3 and h.count() == n
        gl2, gd2 = h.search(q, k)
        np.testing.assert_array_equal(gl, gl2)
        # merge a plain index into the sharded one and the other way round
        Y = rng.standard_normal((50, d)).astype(np.float32)
        plain.add_batch(Y)
        old, new = h.merge(plain, np.arange(0, 50, 2))
        assert list(old) == list(range(0, 50, 2)) and list(new) == list(range(n, n + 25))
        assert np.array_equal(h.get_vector(n + 3), Y[6]) and h.count() == n + 25
        old, new = plain.merge(h, np.array([7, 8, 10 ** 9]))
        assert list(old) == [7, 8] and list(new) == [50, 51]
        assert np.array_equal(plain.get_vector(51), X[8])
    finally:
        h.close()
        plain.close()

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
        with pytest.raises(gpu.LanceError, match="not supported on a sharded handle"):
            h.create_index(8, 4)
        with pytest.raises(gpu.LanceError, match="not supported on a sharded handle"):
            h.save()
        with pytest.raises(gpu.LanceError, match="expected query dimension"):
            h.search(np.zeros(d + 1, np.float32), k)
    finally:
        h.close()
    with pytest.raises(gpu.LanceError, match="does not exist"):
        gpu.LanceHandle.create_sharded([0, 99], d)

"""Extra columns on the device + the pushed-down WHERE evaluated there, through the C ABI.

Checker = the host mirror's evaluator (csrc/host/predicate.cpp, `host.Columns.evaluate`) over the same values, and the
oracle's filtered flat search.  Reference behaviour: src/lance_optimizer.cpp:204-344 (grammar),
test/sql/lance_optimizer_filter.test:23-99 (prefilter semantics), src/lance_index.cpp:321-359 (Arrow ingest)."""
import numpy as np
import pytest

from parity import assert_topk_parity

pytestmark = pytest.mark.gpu


def _check(oracle_mod, X, q, k, metric, got_l, got_d, mask=None, what=""):
    """got vs the oracle's truth ranking under the same allow mask (ids exact up to fp64 ties, distances 1e-4 rel)."""
    allow = None if mask is None else oracle_mod.make_bitmap(mask)
    wl, _, wd64 = oracle_mod.flat_search(X, q, k, metric, allow)
    assert_topk_parity(got_l, got_d, wl, wd64, lambda lab: oracle_mod.distance_f64(metric, q, X[lab]), what=what)


PREDICATES = [
    "score > 20", "score < 1", "score >= 10 AND score <= 30", "(score < 5 OR score > 95)", "NOT (score = 50)",
    "score != 7", "score IN (1, 2, 3, 99)", "score NOT IN (1, 2, 3)", "score IN (1, NULL)", "score NOT IN (1, NULL)",
    "score IS NULL", "score IS NOT NULL", "50 < score", "score > 19.5", "weight > 0.25",
    "weight <= -0.5 AND score > 10", "weight IN (0.0, 1.5)", "lang = 'en'", "lang != 'en'", "lang IN ('en', 'fr')",
    "lang NOT IN ('en', 'fr')", "lang > 'es'", "lang = 'zz'", "lang = 'it''s'", "lang IS NULL", "flag = true",
    "flag = false AND lang = 'es'", "NOT (flag = true)",
    "(lang = 'en' AND score > 20) OR (lang = 'fr' AND weight < 0.0)",
    "NOT ((score > 20 AND flag = true) OR lang IS NULL)", "score = NULL", "lang = 5", "score = 'x'",
    "NOT (score = NULL)", "score > 20 AND score = NULL", "score > 20 OR score = NULL",
]


def _make_columns(rng, n):
    score = [None if rng.random() < 0.05 else int(rng.integers(0, 100)) for _ in range(n)]
    weight = [None if rng.random() < 0.05 else float(np.round(rng.normal(), 2)) for _ in range(n)]
    langs = ["en", "fr", "es", "it's", "de"]
    lang = [None if rng.random() < 0.05 else langs[int(rng.integers(0, len(langs)))] for _ in range(n)]
    flag = [None if rng.random() < 0.05 else bool(rng.integers(0, 2)) for _ in range(n)]
    return {"score": score, "weight": weight, "lang": lang, "flag": flag}


def _push(h, cols, first=0):
    h.set_column("score", np.array([0 if v is None else v for v in cols["score"]], np.int64), first,
                 valid=[v is not None for v in cols["score"]])
    h.set_column("weight", np.array([0.0 if v is None else v for v in cols["weight"]], np.float64), first,
                 valid=[v is not None for v in cols["weight"]])
    h.set_column("lang", cols["lang"], first, type="str")
    h.set_column("flag", np.array([False if v is None else v for v in cols["flag"]], bool), first,
                 valid=[v is not None for v in cols["flag"]])


def _host_columns(host, cols):
    hc = host.Columns()
    for name, vals in cols.items():
        hc.append(name, vals)
    return hc


def _bits(bm, n):
    return np.unpackbits(bm.view(np.uint8), bitorder="little")[:n].astype(bool)


@pytest.mark.parametrize("n", [1, 31, 32, 33, 5000])
def test_device_predicate_matches_host_evaluator(gpu, host, n):
    rng = np.random.default_rng(100 + n)
    cols = _make_columns(rng, n)
    # every column needs one non-null value for the host checker to type it
    cols["score"][0], cols["weight"][0], cols["lang"][0], cols["flag"][0] = 5, 0.5, "en", True
    with gpu.LanceHandle(8) as h:
        h.add_batch(rng.standard_normal((n, 8)).astype(np.float32))
        _push(h, cols)
        hc = _host_columns(host, cols)
        for pred in PREDICATES:
            bm, bits, dptr, _ = h.predicate_bitmap(pred)
            assert bits == n and dptr != 0
            got = _bits(bm, n)
            want = hc.evaluate(pred, n)
            assert np.array_equal(got, want), (pred, np.flatnonzero(got != want)[:8])
            tail = np.unpackbits(bm.view(np.uint8), bitorder="little")[n:]
            assert not tail.any(), pred  # bits beyond the table stay clear


def test_predicate_cache_follows_appends_and_column_writes(gpu):
    rng = np.random.default_rng(5)
    with gpu.LanceHandle(4) as h:
        h.add_batch(rng.standard_normal((100, 4)).astype(np.float32))
        h.set_column("score", np.arange(100, dtype=np.int64))
        bm, bits, _, cached = h.predicate_bitmap("score >= 90")
        assert bits == 100 and not cached and _bits(bm, 100).sum() == 10
        assert h.predicate_bitmap("score >= 90")[3] is True  # second call: served from the device cache
        h.add_batch(rng.standard_normal((50, 4)).astype(np.float32))
        bm, bits, _, cached = h.predicate_bitmap("score >= 90")
        assert bits == 150 and not cached and _bits(bm, 150).sum() == 10  # new rows have no value: NULL, filtered
        h.set_column("score", np.full(50, 95, np.int64), first_label=100)
        bm, bits, _, cached = h.predicate_bitmap("score >= 90")
        assert not cached and _bits(bm, 150).sum() == 60


def test_columns_survive_table_growth(gpu):
    """The vector table doubles (store.cu ensure_capacity): the columns must be carried into the larger buffers."""
    rng = np.random.default_rng(6)
    with gpu.LanceHandle(4) as h:
        total = 0
        for step in range(5):
            m = 3000
            h.add_batch(rng.standard_normal((m, 4)).astype(np.float32))
            h.set_column("score", np.arange(total, total + m, dtype=np.int64), first_label=total)
            total += m
        bm, bits, _, _ = h.predicate_bitmap("score >= 100 AND score < 14000")
        got = _bits(bm, total)
        assert bits == total and np.array_equal(np.flatnonzero(got), np.arange(100, 14000))


def test_malformed_predicates_fail_with_a_message(gpu):
    with gpu.LanceHandle(4) as h:
        h.add_batch(np.zeros((4, 4), np.float32))
        h.set_column("score", np.arange(4, dtype=np.int64))
        for bad in ["nosuch > 1", "score >", "score > 1 AND", "(score > 1", "score ~ 3", "score > 1 garbage"]:
            with pytest.raises(gpu.LanceError, match="lance predicate"):
                h.search_predicate(np.zeros(4, np.float32), 2, bad)
        with pytest.raises(gpu.LanceError, match="fall outside the table"):
            h.set_column("score", np.arange(10, dtype=np.int64))
        with pytest.raises(gpu.LanceError, match="another type"):
            h.set_column("score", np.zeros(4, np.float64))
        with pytest.raises(gpu.LanceError, match="no extra column"):
            h.create_scalar_index("nosuch")
        h.create_scalar_index("score", "BTREE")  # src/lance_index.cpp:481-486: nothing to build, succeeds


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_predicate_search_matches_oracle_prefilter(gpu, oracle_mod, metric):
    """Single queries (fp32 scan) and a batch (tcgen05 path, >= 65536 rows) under a predicate vs the oracle's filtered
    flat search with the same mask."""
    rng = np.random.default_rng(11)
    n, d, k = 70000, 64, 10
    X = rng.standard_normal((n, d)).astype(np.float32)
    score = rng.integers(0, 100, n)
    lang = np.array(["en", "fr", "es"])[rng.integers(0, 3, n)]
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X)
        h.set_column("score", score.astype(np.int64))
        h.set_column("lang", list(lang), type="str")
        for pred, mask in [("score > 20", score > 20), ("score < 1", score < 1),
                           ("lang = 'fr' AND score >= 50", (lang == "fr") & (score >= 50)),
                           ("lang IN ('en', 'es') OR score = 3", np.isin(lang, ["en", "es"]) | (score == 3))]:
            Q = rng.standard_normal((12, d)).astype(np.float32)
            for q in Q[:3]:
                gl, gd = h.search_predicate(q, k, pred)
                _check(oracle_mod, X, q, k, metric, gl, gd, mask, pred)
            L, D, Cn = h.search_batch_predicate(Q, k, pred)
            for i, q in enumerate(Q):
                _check(oracle_mod, X, q, k, metric, L[i, :Cn[i]], D[i, :Cn[i]], mask, f"batch {pred} q{i}")
        gl, gd = h.search_predicate(Q[0], k, None)  # no predicate = no filter, same call
        _check(oracle_mod, X, Q[0], k, metric, gl, gd)


def test_predicate_and_tombstones_compose(gpu, oracle_mod):
    rng = np.random.default_rng(12)
    n, d, k = 3000, 16, 8
    X = rng.standard_normal((n, d)).astype(np.float32)
    score = rng.integers(0, 10, n)
    with gpu.LanceHandle(d) as h:
        h.add_batch(X)
        h.set_column("score", score.astype(np.int64))
        dead = rng.choice(n, 500, replace=False)
        h.delete_batch(dead.astype(np.int64))
        mask = score >= 5
        mask[dead] = False
        q = rng.standard_normal(d).astype(np.float32)
        gl, gd = h.search_predicate(q, k, "score >= 5")
        _check(oracle_mod, X, q, k, "l2", gl, gd, mask)


def test_ivfpq_search_with_predicate(gpu):
    """The prefilter reaches the ADC scan as well (a15 over an IVF_PQ index): same result as the host-bitmap call."""
    rng = np.random.default_rng(13)
    n, d, k = 20000, 32, 10
    X = rng.standard_normal((n, d)).astype(np.float32)
    score = rng.integers(0, 100, n)
    with gpu.LanceHandle(d) as h:
        h.add_batch(X)
        h.set_column("score", score.astype(np.int64))
        h.create_index(16, 8)
        q = rng.standard_normal(d).astype(np.float32)
        gl, gd = h.search_predicate(q, k, "score < 30", nprobes=16)
        assert len(gl) == k and (score[gl] < 30).all()
        al, ad = h.search(q, k, nprobes=16, allow=gpu.make_bitmap(score < 30))
        assert np.array_equal(gl, al) and np.array_equal(gd, ad)


def _arrow_batch(pa, X, extras):
    d = X.shape[1]
    vec = pa.FixedSizeListArray.from_arrays(pa.array(X.reshape(-1), pa.float32()), d)
    return pa.StructArray.from_arrays([vec] + list(extras.values()), ["vector"] + list(extras.keys()))


def _export(sa):
    """pyarrow's C data interface export: (keep-alive handles, schema address, array address)."""
    from pyarrow.cffi import ffi

    c_schema = ffi.new("struct ArrowSchema*")
    c_array = ffi.new("struct ArrowArray*")
    sp, ap = int(ffi.cast("uintptr_t", c_schema)), int(ffi.cast("uintptr_t", c_array))
    sa._export_to_c(ap, sp)
    return (c_schema, c_array, ffi), sp, ap


def test_arrow_ingest_creates_columns_and_filters(gpu, host, oracle_mod):
    """lance_create_detached_from_arrow + lance_detached_add_batch_arrow over pyarrow's C data interface export: the
    path src/lance_index.cpp:286-359 drives for an index with extra columns."""
    pa = pytest.importorskip("pyarrow")
    rng = np.random.default_rng(21)
    n, d, k = 4000, 24, 5
    X = rng.standard_normal((n, d)).astype(np.float32)
    cols = _make_columns(rng, n)
    cols["score"][0], cols["weight"][0], cols["lang"][0], cols["flag"][0] = 5, 0.5, "en", True
    big = [None if v is None else v * (1 << 33) for v in cols["score"]]

    def batch(lo, hi):
        return _arrow_batch(pa, X[lo:hi], {
            "score": pa.array(cols["score"][lo:hi], pa.int32()),  # INTEGER
            "big": pa.array(big[lo:hi], pa.int64()),  # BIGINT
            "weight": pa.array(cols["weight"][lo:hi], pa.float64()),
            "lang": pa.array(cols["lang"][lo:hi], pa.string()),
            "flag": pa.array(cols["flag"][lo:hi], pa.bool_()),
        })

    keep1, sp, ap = _export(batch(0, 1500))
    h = gpu.LanceHandle.create_from_arrow(sp, "l2")
    try:
        assert h.dim == d and h.has_extra_columns()
        labels = h.add_batch_arrow(sp, ap, 1500)
        assert np.array_equal(labels, np.arange(1500))
        assert keep1[1].release == keep1[2].NULL  # consumed (rust_lib/src/ffi.rs:148-182)
        # second batch: a SLICE of a larger array, so every child carries a non-zero offset
        keep2, sp2, ap2 = _export(batch(1000, n).slice(500))
        labels = h.add_batch_arrow(sp2, ap2, n - 1500)
        assert np.array_equal(labels, np.arange(1500, n))
        assert h.count() == n
        hc = _host_columns(host, {"score": cols["score"], "big": big, "weight": cols["weight"], "lang": cols["lang"],
                                  "flag": cols["flag"]})
        for pred in ["score > 20", "big >= 429496729600", "weight < 0.0", "lang = 'en'", "flag = true",
                     "lang IS NULL OR flag = false", "score IN (1, 2, 3) AND lang != 'fr'"]:
            bm, bits, _, _ = h.predicate_bitmap(pred)
            assert bits == n and np.array_equal(_bits(bm, n), hc.evaluate(pred, n)), pred
        q = rng.standard_normal(d).astype(np.float32)
        mask = hc.evaluate("lang = 'en' AND score > 20", n)
        gl, gd = h.search_predicate(q, k, "lang = 'en' AND score > 20")
        _check(oracle_mod, X, q, k, "l2", gl, gd, mask)
    finally:
        h.close()


def test_arrow_ingest_rejects_bad_input(gpu):
    pa = pytest.importorskip("pyarrow")
    keep, sp, ap = _export(pa.StructArray.from_arrays([pa.array([1, 2], pa.int32())], ["score"]))
    with pytest.raises(gpu.LanceError, match="no FixedSizeList column"):
        gpu.LanceHandle.create_from_arrow(sp)
    X = np.zeros((2, 4), np.float32)
    keep, sp, ap = _export(_arrow_batch(pa, X, {"ts": pa.array([1, 2], pa.timestamp("s"))}))
    with pytest.raises(gpu.LanceError, match="unsupported arrow type"):
        gpu.LanceHandle.create_from_arrow(sp)
    keep, sp, ap = _export(_arrow_batch(pa, X, {"score": pa.array([1, 2], pa.int32())}))
    with gpu.LanceHandle(8) as h:  # dimension 8, batch has 4
        with pytest.raises(gpu.LanceError, match="dimension mismatch"):
            h.add_batch_arrow(sp, ap, 2)
        assert keep[1].release == keep[2].NULL  # consumed even on failure, like the Rust side's from_raw


def test_columns_survive_save_open_and_merge(gpu, host, tmp_path):
    rng = np.random.default_rng(31)
    n, d = 1000, 8
    X = rng.standard_normal((n, d)).astype(np.float32)
    cols = _make_columns(rng, n)
    cols["score"][0], cols["weight"][0], cols["lang"][0], cols["flag"][0] = 5, 0.5, "en", True
    hc = _host_columns(host, cols)
    db = str(tmp_path / "db")
    with gpu.LanceHandle(d, "l2", db, "t") as h:
        h.add_batch(X)
        _push(h, cols)
        h.save()
    with gpu.LanceHandle.open(db, "t", "l2") as h2:
        assert h2.count() == n and h2.has_extra_columns()
        for pred in ["score > 20", "lang = 'fr'", "weight < 0.0 OR flag = true", "lang IS NULL"]:
            bm, bits, _, _ = h2.predicate_bitmap(pred)
            assert np.array_equal(_bits(bm, n), hc.evaluate(pred, n)), pred
        # merge the even rows into a handle that already holds rows with another dictionary order
        with gpu.LanceHandle(d) as tgt:
            tgt.add_batch(rng.standard_normal((10, d)).astype(np.float32))
            tgt.set_column("lang", ["de"] * 10, type="str")
            live = np.arange(0, n, 2, dtype=np.int64)
            old, new = tgt.merge(h2, live)
            assert np.array_equal(old, live) and np.array_equal(new, np.arange(10, 10 + len(live)))
            m = 10 + len(live)
            for pred in ["score > 20", "lang = 'fr'", "lang = 'de'", "weight < 0.0", "flag = true", "lang IS NULL"]:
                bm, bits, _, _ = tgt.predicate_bitmap(pred)
                assert np.array_equal(_bits(bm, m)[10:], hc.evaluate(pred, n)[live]), pred
            assert _bits(tgt.predicate_bitmap("lang = 'de'")[0], m)[:10].all()
            assert not _bits(tgt.predicate_bitmap("score IS NOT NULL")[0], m)[:10].any()  # never written: NULL

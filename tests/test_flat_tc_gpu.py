"""GPU parity tests of the batched flat search on the tensor cores (tcgen05 candidate GEMM + exact fp32
re-rank, csrc/flat_gemm.cu), forced with set_flat_path(2), against the oracle and against the fp32 scan.

Same bar as the scan: ids identical except where truth distances tie within 1e-5
relative; distances within 1e-4 relative.  The tensor path is exact by construction (error-bounded candidate
threshold, then fp32 re-rank), so it is also compared id-for-id with the fp32 scan at sizes the oracle would
take too long on."""
import numpy as np
import pytest

from parity import assert_topk_parity

pytestmark = pytest.mark.gpu


def truth_fn(oracle_mod, X, q, metric):
    return lambda lab: oracle_mod.distance_f64(metric, q, X[lab])


def tc_search(h, Q, k, allow=None):
    st0 = h.stats()
    h.set_flat_path(2)
    L, D, C = h.search_batch(Q, k, allow=allow)
    h.set_flat_path(0)
    st1 = h.stats()
    assert st1["flat_tc_batches"] == st0["flat_tc_batches"] + 1, "the tcgen05 path did not run"
    return L, D, C, st1["tc_fallback_queries"] - st0["tc_fallback_queries"]


def check_oracle(oracle_mod, X, Q, L, D, C, k, metric, allow=None, deleted=None, what=""):
    for j in range(Q.shape[0]):
        wl, wd32, wd64 = oracle_mod.flat_search(X, Q[j], k, metric, allow, deleted)
        assert C[j] == len(wl), f"{what} q{j}: count {C[j]} != {len(wl)}"
        assert_topk_parity(L[j, :C[j]], D[j, :C[j]], wl, wd64, truth_fn(oracle_mod, X, Q[j], metric), what=f"{what} q{j}")
        assert (L[j, C[j]:] == -1).all() and np.isinf(D[j, C[j]:]).all()


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
@pytest.mark.parametrize("n,d,k,nq", [(70_000, 128, 10, 130), (66_000, 768, 100, 64), (65_537, 100, 7, 9),
                                      (80_000, 384, 10, 257), (70_001, 16, 3, 33),
                                      # 4 and 6 query tiles on the K-streamed kernel: cluster pairs sharing X tiles by TMA multicast
                                      (70_000, 320, 10, 512), (66_000, 768, 20, 700)])
def test_tc_random_parity(gpu, oracle_mod, metric, n, d, k, nq):
    rng = np.random.default_rng(0xBEEF + n + d)
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X[: n // 3])
        h.add_batch(X[n // 3:])
        L, D, C, fb = tc_search(h, Q, k)
        assert fb == 0, "random data must not overflow candidate lists"
        sel = np.unique(np.concatenate([np.arange(min(nq, 6)), [nq - 1]]))
        check_oracle(oracle_mod, X, Q[sel], L[sel], D[sel], C[sel], k, metric, what=f"tc {metric} n={n} d={d}")
        # every query: identical to the fp32 scan of the same library
        h.set_flat_path(1)
        L1, D1, C1 = h.search_batch(Q, k)
        h.set_flat_path(0)
        np.testing.assert_array_equal(C, C1)
        mism = (L != L1).any(axis=1)
        for j in np.nonzero(mism)[0]:
            # a swap is only legal between (near-)equal distances
            assert_topk_parity(L[j], D[j], L1[j], D1[j].astype(np.float64), truth_fn(oracle_mod, X, Q[j], metric),
                               what=f"tc vs scan q{j}")
        np.testing.assert_allclose(D, D1, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_tc_prefilter_and_deletes(gpu, oracle_mod, metric):
    n, d, k, nq = 90_000, 64, 10, 40
    rng = np.random.default_rng(77)
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    mask = rng.random(n) < 0.6
    dele = rng.choice(n, 5000, replace=False)
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X)
        h.delete_batch(dele)
        allow = gpu.make_bitmap(mask)
        deleted = np.zeros(n, bool)
        deleted[dele] = True
        L, D, C, fb = tc_search(h, Q, k, allow=allow)
        check_oracle(oracle_mod, X, Q[:8], L[:8], D[:8], C[:8], k, metric, allow, oracle_mod.make_bitmap(deleted),
                     what=f"tc filter {metric}")
        for j in range(nq):
            got = L[j, :C[j]]
            assert mask[got].all() and not deleted[got].any()


def test_tc_scaled_data_and_large_norm_spread(gpu, oracle_mod):
    """Rows with very different norms: the error bound uses max |x|, the result must still be exact."""
    n, d, k, nq = 70_000, 128, 10, 16
    rng = np.random.default_rng(5)
    X = rng.standard_normal((n, d), dtype=np.float32) * rng.uniform(0.01, 30.0, (n, 1)).astype(np.float32)
    Q = rng.standard_normal((nq, d), dtype=np.float32) * 7.0
    for metric in ("l2", "dot", "cosine"):
        with gpu.LanceHandle(d, metric) as h:
            h.add_batch(X)
            L, D, C, fb = tc_search(h, Q, k)
            check_oracle(oracle_mod, X, Q[:6], L[:6], D[:6], C[:6], k, metric, what=f"tc scaled {metric}")


def test_tc_degenerate_ties_fall_back_to_exact_scan(gpu, oracle_mod):
    """Masses of identical rows overflow the candidate lists: those queries are re-run by the fp32 scan."""
    n, d, k, nq = 70_000, 32, 5, 8
    rng = np.random.default_rng(9)
    base = rng.standard_normal((1, d), dtype=np.float32)
    X = np.repeat(base, n, axis=0)
    X[:100] = rng.standard_normal((100, d), dtype=np.float32) * 8.0 + 20.0  # far away: the duplicates are nearest
    Q = base + 0.1 * rng.standard_normal((nq, d), dtype=np.float32)
    with gpu.LanceHandle(d, "l2") as h:
        h.add_batch(X)
        L, D, C, fb = tc_search(h, Q, k)
        assert fb > 0
        check_oracle(oracle_mod, X, Q, L, D, C, k, "l2", what="tc ties")


def test_tc_few_live_rows_and_k_larger_than_survivors(gpu, oracle_mod):
    n, d, k, nq = 66_000, 64, 50, 12
    rng = np.random.default_rng(11)
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    mask = np.zeros(n, bool)
    mask[rng.choice(n, 20, replace=False)] = True  # fewer survivors than k
    with gpu.LanceHandle(d, "dot") as h:
        h.add_batch(X)
        L, D, C, fb = tc_search(h, Q, k, allow=gpu.make_bitmap(mask))
        assert (C == 20).all()
        check_oracle(oracle_mod, X, Q[:4], L[:4], D[:4], C[:4], k, "dot", gpu.make_bitmap(mask), what="tc few")


def test_tc_append_after_search_is_visible(gpu, oracle_mod):
    n, d, k = 70_000, 128, 10
    rng = np.random.default_rng(13)
    X = rng.standard_normal((n + 500, d), dtype=np.float32)
    Q = rng.standard_normal((16, d), dtype=np.float32)
    X[n + 7] = Q[3] * 1.0  # will be the nearest neighbour of query 3 once appended
    with gpu.LanceHandle(d, "l2") as h:
        h.add_batch(X[:n])
        tc_search(h, Q, k)
        h.add_batch(X[n:])
        L, D, C, fb = tc_search(h, Q, k)
        assert L[3, 0] == n + 7 and D[3, 0] < 1e-6
        check_oracle(oracle_mod, X, Q[:4], L[:4], D[:4], C[:4], k, "l2", what="tc append")


def test_tc_auto_mode_picks_tensor_path_for_batches(gpu):
    n, d = 70_000, 128
    rng = np.random.default_rng(17)
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((64, d), dtype=np.float32)
    with gpu.LanceHandle(d, "dot") as h:
        h.add_batch(X)
        s0 = h.stats()
        h.search_batch(Q, 10)
        h.search(Q[0], 10)
        s1 = h.stats()
        assert s1["flat_tc_batches"] == s0["flat_tc_batches"] + 1
        assert s1["flat_simt_batches"] == s0["flat_simt_batches"] + 1


def _async_search(h, Q, k):
    """lance_b200_search_device_async + lance_b200_search_finish with device buffers; returns host arrays, the
    overflow word the search wrote on-stream, and the number of queries finish re-ran."""
    import torch

    dev = torch.device("cuda", 0)
    nq = Q.shape[0]
    Qd = torch.from_numpy(Q).to(dev)
    L = torch.full((nq, k), -7, dtype=torch.int64, device=dev)
    D = torch.zeros((nq, k), dtype=torch.float32, device=dev)
    C = torch.zeros(nq, dtype=torch.int32, device=dev)
    O = torch.full((1,), 99, dtype=torch.int32, device=dev)
    sp = torch.cuda.current_stream().cuda_stream
    h.set_flat_path(2)
    h.search_device_async(Qd.data_ptr(), nq, k, L.data_ptr(), D.data_ptr(), C.data_ptr(), stream=sp, d_overflow=O.data_ptr())
    rerun = h.search_finish(sp)
    h.set_flat_path(0)
    return L.cpu().numpy(), D.cpu().numpy(), C.cpu().numpy(), int(O.item()), rerun


def test_tc_async_search_and_finish_match_the_synchronous_call(gpu, oracle_mod):
    n, d, k, nq = 70_000, 128, 10, 96
    rng = np.random.default_rng(21)
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    with gpu.LanceHandle(d, "dot") as h:
        h.add_batch(X)
        L0, D0, C0, _ = tc_search(h, Q, k)
        L, D, C, ovf, rerun = _async_search(h, Q, k)
        assert ovf == 0 and rerun == 0
        assert (L == L0).all() and (C == C0).all() and np.array_equal(D, D0)
        assert h.search_finish(0) == 0  # nothing pending any more


def test_tc_async_overflow_is_reported_and_settled_by_finish(gpu, oracle_mod):
    n, d, k, nq = 70_000, 32, 5, 8
    rng = np.random.default_rng(9)
    base = rng.standard_normal((1, d), dtype=np.float32)
    X = np.repeat(base, n, axis=0)
    X[:100] = rng.standard_normal((100, d), dtype=np.float32) * 8.0 + 20.0
    Q = base + 0.1 * rng.standard_normal((nq, d), dtype=np.float32)
    with gpu.LanceHandle(d, "l2") as h:
        h.add_batch(X)
        L, D, C, ovf, rerun = _async_search(h, Q, k)
        assert ovf > 0 and rerun == ovf
        check_oracle(oracle_mod, X, Q, L, D, C, k, "l2", what="tc async ties")


def test_full_size_c2_tensor_path_equals_scan(gpu):
    """BASELINE.json config 2 at its full size (1M x 768, cosine, k=100), rows generated on the device: the tcgen05
    batch path and the fp32 streaming scan (two independent kernels) must return the same lists; size-independent
    properties: sorted ascending, no duplicate labels, k results per query."""
    import torch

    n, d, k, nq = 1_000_000, 768, 100, 24
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev)
    g.manual_seed(123)
    with gpu.LanceHandle(d, "cosine") as h:
        h.reserve(n)
        sp = torch.cuda.current_stream().cuda_stream
        for c0 in range(0, n, 250_000):
            blk = torch.randn((250_000, d), generator=g, device=dev, dtype=torch.float32)
            h.add_batch_device(blk.data_ptr(), 250_000, d, sp)
            torch.cuda.synchronize()
        Q = torch.randn((nq, d), generator=g, device=dev, dtype=torch.float32).cpu().numpy()
        L, D, C, fb = tc_search(h, Q, k)
        assert fb == 0 and (C == k).all()
        assert (np.diff(D, axis=1) >= 0).all()
        for j in range(nq):
            assert len(set(L[j].tolist())) == k
        for j in range(0, nq, 5):
            l1, d1 = h.search(Q[j], k)  # single query -> flat_scan_kernel
            # same ids except where neighbours tie within 1e-5 relative; distances within 1e-4 relative
            if list(l1) != list(L[j]):
                assert sorted(l1) == sorted(L[j]) or np.allclose(np.sort(d1), np.sort(D[j]), rtol=1e-5), f"q{j}"
            np.testing.assert_allclose(D[j], d1, rtol=1e-4, atol=1e-6)


def test_sharded_step_on_one_gpu_packed_record_and_strided_merge(gpu, oracle_mod):
    """The multi-GPU step (sharding.ShardedSearch: async search into the packed record -> gather -> strided device
    merge -> settle) with world = 1, and the strided merge over two hand-made shards of the same table."""
    import importlib

    import torch

    sharding = importlib.import_module("duckdb_lancedb_b200.sharding")
    n, d, k, nq = 70_000, 64, 10, 40
    rng = np.random.default_rng(31)
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    dev = torch.device("cuda", 0)
    Qd = torch.from_numpy(Q).to(dev)
    sp = torch.cuda.current_stream().cuda_stream
    with gpu.LanceHandle(d, "l2") as h:
        h.add_batch(X)
        ss = sharding.ShardedSearch(gpu, h, 1, nq, k, dev, 0)
        ML, MD, MC = ss.step(Qd.data_ptr(), sp)
        assert ss.settle() == 0
        L, D, C = ML.cpu().numpy(), MD.cpu().numpy(), MC.cpu().numpy()
        check_oracle(oracle_mod, X, Q[:6], L[:6], D[:6], C[:6], k, "l2", what="sharded world=1")
        L1, D1, C1 = h.search_batch(Q, k)
        assert (L == L1).all() and (C == C1).all()
    # two shards (label ranges) searched one after the other into a 2-rank record buffer, merged by the strided kernel
    half = n // 2
    g2 = sharding.ShardGather(2, nq, k, dev)
    for r, (lo, hi) in enumerate(((0, half), (half, n))):
        with gpu.LanceHandle(d, "l2") as h:
            h.set_label_base(lo)
            h.add_batch(X[lo:hi])
            h.search_device(Qd.data_ptr(), nq, k, g2.L.data_ptr(), g2.D.data_ptr(), g2.C.data_ptr(), stream=sp)
            torch.cuda.synchronize()
            g2.all[r].copy_(g2.rec)
    gpu.merge_topk_strided_device(0, g2.GL.data_ptr(), g2.GD.data_ptr(), g2.GC.data_ptr(), g2.GL.stride(0), g2.GD.stride(0),
                                  g2.GC.stride(0), 2, nq, k, g2.ML.data_ptr(), g2.MD.data_ptr(), g2.MC.data_ptr(), stream=sp)
    torch.cuda.synchronize()
    assert (g2.ML.cpu().numpy() == L1).all() and (g2.MC.cpu().numpy() == C1).all()
    np.testing.assert_allclose(g2.MD.cpu().numpy(), D1, rtol=1e-5)


def test_tc_concurrent_batched_searches_right_after_an_append(gpu, oracle_mod):
    """Two threads issue their first batched searches (each on its own library stream) right after add_batch: the
    bf16 shadow and max|x| are built on ONE of the streams, the other must wait for them (cudaEvent) instead of
    reading an unwritten shadow."""
    import threading

    n, d, k, nq = 300_000, 128, 10, 64
    rng = np.random.default_rng(41)
    X = rng.standard_normal((n, d), dtype=np.float32)
    Qs = [rng.standard_normal((nq, d), dtype=np.float32) for _ in range(4)]
    for rep in range(3):
        with gpu.LanceHandle(d, "dot") as h:
            h.add_batch(X)
            out = [None] * len(Qs)
            start = threading.Barrier(len(Qs))

            def work(i):
                start.wait()
                out[i] = h.search_batch(Qs[i], k)

            ts = [threading.Thread(target=work, args=(i,)) for i in range(len(Qs))]
            [t.start() for t in ts]
            [t.join() for t in ts]
            assert h.stats()["flat_tc_batches"] == len(Qs)
            for i, (L, D, C) in enumerate(out):
                check_oracle(oracle_mod, X, Qs[i][:3], L[:3], D[:3], C[:3], k, "dot", what=f"concurrent rep{rep} t{i}")
            # all threads' full batches equal a quiet re-run
            for i, (L, D, C) in enumerate(out):
                L2, D2, C2 = h.search_batch(Qs[i], k)
                assert (L == L2).all() and (C == C2).all()


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_tc_nan_and_inf_rows_do_not_poison_the_batch(gpu, oracle_mod, metric):
    """A NaN row (and, for l2, an Inf row): max|x| (the error bound of every query) must stay finite; with enough finite rows
    the results are those of the table without the two rows, and the batched path equals the single-query scan."""
    n, d, k, nq = 70_000, 64, 10, 16
    rng = np.random.default_rng(43)
    X = rng.standard_normal((n, d), dtype=np.float32)
    X[1234, 5] = np.nan
    if metric == "l2":  # an infinite component is infinitely FAR under l2 (under dot it would be infinitely near)
        X[4321, 7] = np.inf
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X)
        L, D, C, fb = tc_search(h, Q, k)
        assert (C == k).all() and np.isfinite(D).all()
        assert not np.isin(L, [1234, 4321]).any()
        dele = oracle_mod.make_bitmap(np.isin(np.arange(n), [1234, 4321]))
        check_oracle(oracle_mod, X, Q[:4], L[:4], D[:4], C[:4], k, metric, None, dele, what=f"nan rows {metric}")
        # fewer finite survivors than k.  l2 / dot: the exact scan takes over for those queries and ranks the NaN row
        # last, like the oracle; cosine: a row without a usable norm has distance 1 in every exact kernel (and in the
        # oracle), so the tensor path serves it as a zero row
        mask = np.zeros(n, bool)
        mask[[10, 20, 30, 1234]] = True
        bm = gpu.make_bitmap(mask)
        L2, D2, C2, fb2 = tc_search(h, Q, k, allow=bm)
        assert (C2 == 4).all() and fb2 == (0 if metric == "cosine" else nq)
        check_oracle(oracle_mod, X, Q, L2, D2, C2, k, metric, bm, what=f"nan row among 4 survivors {metric}")
        for j in range(nq):
            l1, d1 = h.search(Q[j], k, allow=bm)
            assert list(l1) == list(L2[j, :4])
            if metric != "cosine":
                assert l1[3] == 1234 and np.isnan(d1[3])


def test_tc_nonfinite_query_goes_to_the_exact_scan(gpu, oracle_mod):
    n, d, k, nq = 70_000, 32, 5, 9
    rng = np.random.default_rng(47)
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    Q[3, 0] = np.inf
    with gpu.LanceHandle(d, "dot") as h:
        h.add_batch(X)
        L, D, C, fb = tc_search(h, Q, k)
        assert fb == 1
        ok = [j for j in range(nq) if j != 3]
        check_oracle(oracle_mod, X, Q[ok][:4], L[ok][:4], D[ok][:4], C[ok][:4], k, "dot", what="inf query batch")
        l1, d1 = h.search(Q[3], k)
        assert list(l1) == list(L[3, :C[3]])


@pytest.mark.parametrize("metric,d,k,nq", [("cosine", 128, 10, 64), ("dot", 128, 10, 256), ("cosine", 768, 100, 32)])
def test_tc_clustered_unit_norm_embeddings_vs_oracle(gpu, oracle_mod, metric, d, k, nq):
    """Clustered, L2-normalised data (bench.py --data clustered: a mixture of 1000 Gaussians, sigma 0.05): the k-th score
    gap is far smaller than on N(0,1), which is what stresses the bf16 candidate bound and the candidate lists.  The
    result must still be the exact search; queries the bound cannot certify are re-run (batched) by the fp32 scan."""
    import torch

    import bench

    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev)
    g.manual_seed(123)
    n = 150_000
    X_t = bench._clustered(n, d, g, dev)
    Q_t = bench._clustered(nq, d, g, dev)  # queries from the same mixture
    X, Q = X_t.cpu().numpy(), Q_t.cpu().numpy()
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X)
        L, D, C, fb = tc_search(h, Q, k)
        st = h.stats()
        assert st["flat_tc_batches"] >= 1
        assert fb <= nq // 8, f"{fb} of {nq} queries fell back to the exact scan on clustered data"
        check_oracle(oracle_mod, X, Q[:8], L[:8], D[:8], C[:8], k, metric, what=f"tc clustered {metric} d={d}")


@pytest.mark.parametrize("metric,d,k,seed_tiles", [("dot", 128, 10, 64), ("l2", 64, 10, 64), ("cosine", 200, 10, 64),
                                                   # K-streamed kernel (half-tile maxima), plain and with the l2 bias epilogue
                                                   ("cosine", 320, 10, 64), ("l2", 704, 10, 64),
                                                   # larger k (the k-th of 128 / 256 maxima), odd dimension
                                                   ("l2", 100, 50, 128), ("cosine", 320, 100, 128)])
def test_tc_tile_max_seeding_pass(gpu, oracle_mod, metric, d, k, seed_tiles, monkeypatch):
    """The threshold seeding pass of the tensor-memory kernels (flat_gemm.cu "Seeding": per-tile maxima, no emission),
    forced at a size the oracle can check (64 seed tiles instead of 1024), with deletions inside and outside the seeded
    rows, a prefilter, a partial last tile and rows sorted so that the best rows of some queries share ONE tile (the
    case in which the k-th tile maximum is a loose bound: still a bound, the result must not change)."""
    n, nq = 150_001, 300
    rng = np.random.default_rng(4242 + d)
    monkeypatch.setenv("LANCE_B200_TC_GROWTH", "2")  # the pass needs 2 x growth x seed tiles of table: keep it small here
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    # tile 3 (rows 384..511) holds near-copies of query 0: its best rows sit in one tile.  (Looser copies for the larger k:
    # a cosine distance of 1e-3 is 1 - 0.999 in fp32, which cannot rank 100 such rows to 1e-5 relative — nor can the reference.)
    noise = 0.05 if k <= 10 else 0.3
    X[384:512] = Q[0] * (1.0 + 0.01 * rng.standard_normal((128, 1), dtype=np.float32)) + \
        noise * rng.standard_normal((128, d), dtype=np.float32)
    mask = rng.random(n) < 0.7
    dele = np.concatenate([rng.choice(8192, 600, replace=False), 8192 + rng.choice(n - 8192, 3000, replace=False)])
    deleted = np.zeros(n, bool)
    deleted[dele] = True
    monkeypatch.setenv("LANCE_B200_TC_SEEDMAX", str(seed_tiles))
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X)
        l0 = h.stats()["kernel_launches"]
        L, D, C, fb = tc_search(h, Q, k)
        seeded_launches = h.stats()["kernel_launches"] - l0
        assert fb == 0
        sel = np.array([0, 1, 2, 127, 128, 255, 299])
        check_oracle(oracle_mod, X, Q[sel], L[sel], D[sel], C[sel], k, metric, what=f"seeded {metric}")
        monkeypatch.setenv("LANCE_B200_TC_SEEDMAX", "0")
        l0 = h.stats()["kernel_launches"]
        L0, D0, C0, _ = tc_search(h, Q, k)
        assert h.stats()["kernel_launches"] - l0 != seeded_launches, "the seeding pass did not run"
        np.testing.assert_array_equal(L, L0)
        np.testing.assert_array_equal(D, D0)
        # with tombstones and a prefilter
        monkeypatch.setenv("LANCE_B200_TC_SEEDMAX", str(seed_tiles))
        h.delete_batch(dele)
        allow = gpu.make_bitmap(mask)
        L, D, C, fb = tc_search(h, Q, k, allow=allow)
        check_oracle(oracle_mod, X, Q[sel], L[sel], D[sel], C[sel], k, metric, allow, oracle_mod.make_bitmap(deleted),
                     what=f"seeded filtered {metric}")
        for j in range(nq):
            got = L[j, :C[j]]
            assert mask[got].all() and not deleted[got].any()

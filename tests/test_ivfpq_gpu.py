"""GPU parity tests of the IVF_PQ path (probe -> LUT -> ADC scan -> refine) against the oracle, through the C ABI.

BASELINE.json's criterion: with the same trained index, probed partitions and candidate ids identical, ADC
distances within 1e-4 relative; and (refine_factor=1, src/include/lance_index.hpp:92) the exact post-refine
distances.  "Same trained index" is met both ways: a checker-trained index staged into HBM
(lance_b200_set_ivfpq), and a GPU-trained index exported to the checker (lance_b200_get_ivfpq)."""
import numpy as np
import pytest

from parity import assert_topk_parity

pytestmark = pytest.mark.gpu

ADC_RTOL = 1e-4


def compare_with_oracle(oracle_mod, h, ix, X, q, k, nprobes, rf, metric, allow=None, deleted=None, what=""):
    got = h.ivfpq_search_debug(q, k, nprobes, rf, allow=allow)
    want = ix.search(X, q, k, nprobes, rf, allow, deleted)
    # probed partitions: identical, in order (swaps only between centroids tying within 1e-5 relative)
    wp, wpd = ix.probe(q, min(nprobes, ix.P))
    if list(got["parts"]) != list(wp):
        assert sorted(got["parts"]) == sorted(wp), f"{what}: probed set differs"
    # pre-refine ADC candidates
    adc_of = dict(zip(want["cand_labels"].tolist(), want["cand_adc"].tolist()))
    scale = float(np.abs(want["cand_adc"]).max()) if len(want["cand_adc"]) else 1.0

    def adc_truth(lab):
        if lab in adc_of:
            return adc_of[lab]
        if lab >= ix.n_indexed:  # unindexed tail row: exact distance
            return oracle_mod.distance_f64(metric, q, X[lab])
        # a candidate the oracle cut: recompute its ADC distance with the checker
        pos = int(np.nonzero(ix.row_ids == lab)[0][0])
        p = int(np.searchsorted(ix.part_offsets, pos, side="right") - 1)
        return _adc(oracle_mod, ix, q, p, pos)

    assert len(got["cand_labels"]) == len(want["cand_labels"]), f"{what}: candidate count"
    assert_topk_parity(got["cand_labels"], got["cand_adc"], want["cand_labels"], want["cand_adc"], adc_truth,
                       scale=scale, what=f"{what} ADC candidates")
    # final list
    truth = lambda lab: oracle_mod.distance_f64(metric, q, X[lab])
    if rf >= 1:
        assert_topk_parity(got["labels"], got["dist"], want["labels"], want["dist"], truth, what=f"{what} final")
    else:
        assert_topk_parity(got["labels"], got["dist"], want["labels"], want["dist"], adc_truth, scale=scale,
                           what=f"{what} final (no refine)")
    # the plain reference call returns the same final list
    l2, d2 = h.search(q, k, nprobes=nprobes, refine_factor=rf, allow=allow)
    assert list(l2) == list(got["labels"])
    np.testing.assert_array_equal(d2, got["dist"])
    return got


def _adc(oracle_mod, ix, q, part, pos):
    lut = ix.lut(q, part)
    s = float(lut[np.arange(ix.M), ix.codes[pos]].astype(np.float64).sum())
    return {0: s, 1: 0.5 * s, 2: 1.0 + s}[ix.metric]


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
@pytest.mark.parametrize("d,P,M", [(32, 16, 8), (64, 12, 16), (12, 5, 3), (96, 8, 96)])
def test_checker_trained_index_on_gpu(gpu, oracle_mod, metric, d, P, M):
    rng = np.random.default_rng(100 + d + P)
    n = 4000
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((3, d), dtype=np.float32)
    ix = oracle_mod.IvfPqIndex.train(X, metric, P, M, iters=4)
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X)
        h.set_ivfpq(ix.centroids, ix.codebook, ix.part_offsets, ix.row_ids, ix.codes)
        assert h.stats()["has_index"] == 1
        for j in range(3):
            for nprobes, rf, k in ((4, 1, 10), (P, 3, 10), (1, 0, 5), (3, 2, 37), (P + 5, 1, 10)):
                compare_with_oracle(oracle_mod, h, ix, X, Q[j], k, nprobes, rf, metric,
                                    what=f"{metric} d={d} P={P} M={M} q{j} np={nprobes} rf={rf} k={k}")
        assert h.stats()["ivfpq_batches"] > 0
        # round trip of the staged index
        back = h.get_ivfpq()
        assert back["P"] == P and back["M"] == M and back["n_indexed"] == n
        np.testing.assert_array_equal(back["codes"], ix.codes)
        np.testing.assert_array_equal(back["row_ids"], ix.row_ids)
        np.testing.assert_array_equal(back["centroids"], ix.centroids)
        h.drop_index()
        l, _ = h.search(Q[0], 5)
        wl, _, _ = oracle_mod.flat_search(X, Q[0], 5, metric)
        assert list(l) == list(wl)


@pytest.mark.parametrize("d,M", [(64, 32), (128, 64), (96, 96), (128, 128), (160, 160), (192, 192)])
def test_rotated_scan_layout_all_widths(gpu, oracle_mod, d, M):
    """M % 32 == 0 selects the lane-rotated code layout with the transposed table (adc_scan_kernel<32, M/32>:
    software-pipelined for M <= 96, plain loop above); every width must give the oracle's candidates, with ragged
    lists (sizes not multiples of 32), deletes and a prefilter."""
    rng = np.random.default_rng(900 + M)
    n, P = 5000, 7
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((2, d), dtype=np.float32)
    for metric in ("l2", "dot"):
        ix = oracle_mod.IvfPqIndex.train(X, metric, P, M, iters=3)
        with gpu.LanceHandle(d, metric) as h:
            h.add_batch(X)
            h.set_ivfpq(ix.centroids, ix.codebook, ix.part_offsets, ix.row_ids, ix.codes)
            dead = rng.choice(n, 300, replace=False)
            h.delete_batch(dead)
            deleted = oracle_mod.make_bitmap(np.isin(np.arange(n), dead))
            allow = gpu.make_bitmap(rng.random(n) < 0.3)
            for j in range(2):
                compare_with_oracle(oracle_mod, h, ix, X, Q[j], 10, 3, 1, metric, None, deleted, what=f"rot {metric} M={M} q{j}")
                compare_with_oracle(oracle_mod, h, ix, X, Q[j], 25, P, 2, metric, allow, deleted, what=f"rot {metric} M={M} q{j} allow")
            # a batch (one CTA per query walks all its lists) equals the single-query launches (lists split over CTAs)
            Qb = rng.standard_normal((70, d), dtype=np.float32)
            L, D, C = h.search_batch(Qb, 10, nprobes=4, refine_factor=0)
            for j in (0, 33, 69):
                l, dd = h.search(Qb[j], 10, nprobes=4, refine_factor=0)
                assert list(l) == list(L[j, :C[j]])
                # the (query, partition) term comes from the tiled probe in a batch, from the streaming one otherwise
                np.testing.assert_allclose(dd, D[j, :C[j]], rtol=1e-5)
            back = h.get_ivfpq()
            np.testing.assert_array_equal(back["codes"], ix.codes)  # the exchange layout is not rotated


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_gpu_trained_index_exported_to_checker(gpu, oracle_mod, metric):
    rng = np.random.default_rng(77)
    n, d, P, M = 20_000, 64, 32, 16
    centers = rng.standard_normal((P, d), dtype=np.float32) * 3
    X = (centers[rng.integers(0, P, n)] + rng.standard_normal((n, d), dtype=np.float32)).astype(np.float32)
    Q = (centers[rng.integers(0, P, 4)] + rng.standard_normal((4, d), dtype=np.float32)).astype(np.float32)
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X)
        h.create_index(P, M)  # lance_detached_create_index -> GPU k-means + PQ training
        g = h.get_ivfpq()
        assert (g["P"], g["M"], g["n_indexed"], g["dim"]) == (P, M, n, d)
        off = g["part_offsets"]
        assert off[0] == 0 and off[-1] == n and (np.diff(off) >= 0).all()
        assert sorted(g["row_ids"].tolist()) == list(range(n))  # a permutation of the labels
        for p in range(P):  # ascending labels inside a partition
            seg = g["row_ids"][off[p]:off[p + 1]]
            assert (np.diff(seg) > 0).all()
        assert (np.diff(off) > 0).sum() >= P // 2, "k-means collapsed"
        ix = oracle_mod.IvfPqIndex(metric, d, P, M, g["centroids"], g["codebook"], off, g["row_ids"], g["codes"])
        # the exported assignment is the nearest centroid (spot check) and the codes are the nearest codewords
        Xn = X / np.linalg.norm(X, axis=1, keepdims=True) if metric == "cosine" else X
        for pos in rng.integers(0, n, 50):
            row = int(g["row_ids"][pos])
            p = int(np.searchsorted(off, pos, side="right") - 1)
            if metric == "dot":
                sc = -(g["centroids"].astype(np.float64) @ Xn[row].astype(np.float64))
            else:
                sc = ((g["centroids"].astype(np.float64) - Xn[row].astype(np.float64)) ** 2).sum(1)
            assert sc[p] <= sc.min() + 1e-5 * abs(sc.min()) + 1e-5
        for j in range(4):
            for nprobes, rf in ((8, 1), (32, 2), (2, 0)):
                compare_with_oracle(oracle_mod, h, ix, X, Q[j], 10, nprobes, rf, metric,
                                    what=f"gpu-trained {metric} q{j} np={nprobes} rf={rf}")
        # recall sanity: with all partitions probed and a generous refine the exact answer comes back
        l, dd = h.search(Q[0], 10, nprobes=P, refine_factor=50)
        wl, wd, _ = oracle_mod.flat_search(X, Q[0], 10, metric)
        assert len(set(l.tolist()) & set(wl.tolist())) >= 9


def test_index_with_appends_deletes_and_prefilter(gpu, oracle_mod):
    """Rows appended after the index was built are scanned exactly and merged (lance_basic.test:45-53 semantics
    under an ANN index); deletes and the prefilter apply to both parts (lance_manager.rs:461-471)."""
    rng = np.random.default_rng(31)
    n0, n1, d, P, M = 6000, 1500, 32, 16, 8
    X = rng.standard_normal((n0 + n1, d), dtype=np.float32)
    Q = rng.standard_normal((3, d), dtype=np.float32)
    ix = oracle_mod.IvfPqIndex.train(X[:n0], "l2", P, M, iters=4)
    with gpu.LanceHandle(d, "l2") as h:
        h.add_batch(X[:n0])
        h.set_ivfpq(ix.centroids, ix.codebook, ix.part_offsets, ix.row_ids, ix.codes)
        h.add_batch(X[n0:])
        dead = rng.choice(n0 + n1, 700, replace=False)
        h.delete_batch(dead)
        deleted = oracle_mod.make_bitmap(np.isin(np.arange(n0 + n1), dead))
        for sel in (None, 0.5, 0.02):
            allow = None if sel is None else gpu.make_bitmap(rng.random(n0 + n1) < sel)
            for j in range(3):
                compare_with_oracle(oracle_mod, h, ix, X, Q[j], 10, 6, 1, "l2", allow, deleted, what=f"tail sel={sel} q{j}")
                compare_with_oracle(oracle_mod, h, ix, X, Q[j], 10, 6, 0, "l2", allow, deleted, what=f"tail rf0 sel={sel} q{j}")


def test_batch_equals_singles_and_defaults(gpu, oracle_mod):
    rng = np.random.default_rng(5)
    n, d = 10_000, 48
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((37, d), dtype=np.float32)
    with gpu.LanceHandle(d, "l2") as h:
        h.add_batch(X)
        h.create_index(0, 0)  # defaults: partitions = round(sqrt(n)) = 100, sub-vectors = d/16 = 3
        g = h.get_ivfpq()
        assert g["P"] == 100 and g["M"] == 3
        L, D, C = h.search_batch(Q, 10, nprobes=20, refine_factor=1)
        for j in range(0, 37, 6):
            l, dd = h.search(Q[j], 10, nprobes=20, refine_factor=1)
            assert list(l) == list(L[j, :C[j]])
            np.testing.assert_array_equal(dd, D[j, :C[j]])
        with pytest.raises(gpu.LanceError, match="not divisible"):
            h.create_index(10, 7)
    with gpu.LanceHandle(d, "l2") as h:
        with pytest.raises(gpu.LanceError, match="empty table"):
            h.create_index(4, 4)


@pytest.mark.parametrize("metric", ["cosine", "dot", "l2"])
def test_batched_probe_equals_single_query_probe(gpu, oracle_mod, metric):
    """Batches (>= 8 queries) probe with the tiled contraction, single queries with the streaming scan: the probed
    partitions, and with them the final lists, must agree (and agree with the oracle for a sample of queries)."""
    rng = np.random.default_rng(77)
    n, d, P, M = 30_000, 64, 64, 16
    X = rng.standard_normal((n, d), dtype=np.float32)
    Q = rng.standard_normal((41, d), dtype=np.float32)
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch(X)
        h.create_index(P, M)
        g = h.get_ivfpq()
        ix = oracle_mod.IvfPqIndex(metric, d, g["P"], g["M"], g["centroids"], g["codebook"], g["part_offsets"], g["row_ids"],
                                   g["codes"], n_indexed=g["n_indexed"])
        L, D, C = h.search_batch(Q, 10, nprobes=8, refine_factor=1)
        for j in range(41):
            l, dd = h.search(Q[j], 10, nprobes=8, refine_factor=1)
            assert list(l) == list(L[j, :C[j]]), f"{metric} q{j}"
            np.testing.assert_allclose(dd, D[j, :C[j]], rtol=1e-6)
        for j in (0, 13, 40):
            compare_with_oracle(oracle_mod, h, ix, X, Q[j], 10, 8, 1, metric, what=f"{metric} batched-probe q{j}")


def test_default_sub_vectors_for_awkward_dimensions_and_bad_exchange_rows(gpu, oracle_mod):
    """dim % 8 != 0: lancedb's default would be ONE sub-vector of `dim` floats, which the device scan's lookup-table
    kernel cannot hold for large dims; the default then falls to the widest sub-vector that fits, and an explicit
    infeasible request is rejected at create time (not at the first search)."""
    rng = np.random.default_rng(3)
    n, d = 3000, 300
    X = rng.standard_normal((n, d), dtype=np.float32)
    with gpu.LanceHandle(d, "l2") as h:
        h.add_batch(X)
        h.create_index(8, 0)
        g = h.get_ivfpq()
        assert d % g["M"] == 0 and d // g["M"] <= 200 and g["M"] > 1
        l, dd = h.search(X[17], 5, nprobes=8, refine_factor=4)
        assert l[0] == 17 and dd[0] < 1e-6
        with pytest.raises(gpu.LanceError, match="too large"):
            h.create_index(8, 1)
    # set_ivfpq: row ids that are not a permutation of the indexed prefix are rejected
    n, d, P, M = 2000, 16, 4, 4
    X = rng.standard_normal((n, d), dtype=np.float32)
    ix = oracle_mod.IvfPqIndex.train(X, "l2", P, M, iters=2)
    with gpu.LanceHandle(d, "l2") as h:
        h.add_batch(X)
        bad = ix.row_ids.copy()
        bad[5] = bad[6]
        with pytest.raises(gpu.LanceError, match="permutation"):
            h.set_ivfpq(ix.centroids, ix.codebook, ix.part_offsets, bad, ix.codes)
        h.set_ivfpq(ix.centroids, ix.codebook, ix.part_offsets, ix.row_ids, ix.codes)


def test_indexed_table_survives_save_and_open(gpu, oracle_mod, tmp_path):
    """The reference keeps its ANN index across CHECKPOINT + restart (src/lance_index.cpp:492-587 reopens the Lance
    dataset with its _indices; test/sql/lance_persist.test:20-34): an indexed table saved by lance_b200_save reopens
    INDEXED through lance_open_detached — same partitions, same candidates, same lists — with its deletes, and rows
    inserted after the restart are found through the unindexed tail."""
    import os

    rng = np.random.default_rng(19)
    n, d, P, M = 6000, 32, 16, 8
    X = rng.standard_normal((n + 50, d), dtype=np.float32)
    Q = rng.standard_normal((3, d), dtype=np.float32)
    dead = rng.choice(n, 200, replace=False)
    with gpu.LanceHandle(d, "l2", db_path=str(tmp_path), table_name="t_ann") as h:
        h.add_batch(X[:n])
        h.create_index(P, M)
        h.delete_batch(dead)
        g0 = h.get_ivfpq()
        want = [h.ivfpq_search_debug(Q[j], 10, 6, 1) for j in range(3)]
        h.save()
    with gpu.LanceHandle.open(str(tmp_path), "t_ann", "l2") as h2:
        assert h2.stats()["has_index"] == 1 and h2.count() == n - 200
        g1 = h2.get_ivfpq()
        for key in ("centroids", "codebook", "part_offsets", "row_ids", "codes"):
            np.testing.assert_array_equal(g0[key], g1[key])
        ix = oracle_mod.IvfPqIndex("l2", d, P, M, g1["centroids"], g1["codebook"], g1["part_offsets"], g1["row_ids"], g1["codes"])
        deleted = oracle_mod.make_bitmap(np.isin(np.arange(n + 50), dead))
        for j in range(3):
            got = h2.ivfpq_search_debug(Q[j], 10, 6, 1)
            for key in ("parts", "cand_labels", "labels"):
                assert list(got[key]) == list(want[j][key]), key
            np.testing.assert_array_equal(got["dist"], want[j]["dist"])
            compare_with_oracle(oracle_mod, h2, ix, X[:n], Q[j], 10, 6, 1, "l2", None, deleted, what=f"reopened q{j}")
        # insert after restart: labels continue, the new rows are searched exactly (lance_persist.test:29-34)
        new = h2.add_batch(X[n:])
        assert list(new) == list(range(n, n + 50))
        l, dd = h2.search(X[n + 7], 5, nprobes=6, refine_factor=1)
        assert l[0] == n + 7 and dd[0] < 1e-6
    # a truncated file is an error, not a short table
    path = os.path.join(str(tmp_path), "t_ann.b200", "store.bin")
    size = os.path.getsize(path)
    with open(path, "r+b") as f:
        f.truncate(size - 1000)
    with pytest.raises(gpu.LanceError, match="truncated"):
        gpu.LanceHandle.open(str(tmp_path), "t_ann", "l2")
    with open(path, "r+b") as f:
        f.truncate(200)
    with pytest.raises(gpu.LanceError, match="truncated"):
        gpu.LanceHandle.open(str(tmp_path), "t_ann", "l2")


def test_ivfpq_sharded_by_partition(gpu):
    """whole inverted lists per shard, balanced by length.  Every shard probes the SAME partitions and scans
    the ones it owns; local top-k (ADC -> refine) -> merge by exact distance, the north_star's recipe.  Guarantees checked:
    a one-shard split reproduces the one-device result bit for bit (explicit labels + row-space index exchange); with
    several shards every returned row lies in a probed partition and the merged list is rank by rank at least as near as
    the one-device list (each shard refines ITS k best ADC candidates, a superset of the one-device candidate set).
    The shards sit on device 0 here; bench.py --workload c3 --gpus N puts one on every GPU."""
    import importlib

    import torch

    sharding = importlib.import_module("duckdb_lancedb_b200.sharding")
    g = torch.Generator(device="cuda:0")
    g.manual_seed(5)
    n, d, k, nq, nprobes = 60000, 64, 10, 24, 8
    X = torch.randn((n, d), generator=g, device="cuda:0")
    Q = torch.randn((nq, d), generator=g, device="cuda:0").cpu().numpy()
    with gpu.LanceHandle(d, "l2") as full:
        full.add_batch_device(X.data_ptr(), n, d)
        full.create_index(64, 8)
        index = full.get_ivfpq()
        sizes = np.diff(index["part_offsets"])
        part_of_label = np.empty(n, np.int64)
        for p in range(index["P"]):
            part_of_label[index["row_ids"][index["part_offsets"][p]:index["part_offsets"][p + 1]]] = p
        wantL, wantD, wantC = full.search_batch(Q, k, nprobes=nprobes)
        want1 = [full.search(q, k, nprobes=nprobes) for q in Q[:3]]
        probed = [set(full.ivfpq_search_debug(q, k, nprobes, 1)["parts"].tolist()) for q in Q]
        for world in (1, 3):
            shards = [sharding.ivfpq_partition_shard(gpu, index, X, r, world, "l2")[0] for r in range(world)]
            try:
                owner = sharding.assign_partitions(sizes, world)
                loads = np.array([sizes[owner == r].sum() for r in range(world)])
                assert loads.sum() == n and loads.max() - loads.min() <= sizes.max()
                assert [s.count() for s in shards] == loads.tolist()
                res = [s.search_batch(Q, k, nprobes=nprobes) for s in shards]
                for i in range(nq):  # merge under (distance, label)
                    cand = sorted((float(r[1][i, j]), int(r[0][i, j])) for r in res for j in range(r[2][i]))[:k]
                    gl = [c[1] for c in cand]
                    gd = np.array([c[0] for c in cand], np.float32)
                    if world == 1:
                        assert gl == wantL[i, :wantC[i]].tolist() and np.array_equal(gd, wantD[i, :wantC[i]]), f"q{i}"
                    else:
                        assert len(gl) == wantC[i] and len(set(gl)) == len(gl)
                        assert all(int(part_of_label[l]) in probed[i] for l in gl), f"q{i}: a row outside the probed lists"
                        assert (gd <= wantD[i, :wantC[i]] * (1 + 1e-6)).all(), f"q{i}: farther than the one-device list"
                for q, (wl, wd) in zip(Q[:3], want1):  # the single-query kernel (ivfpq1) on the shards
                    cand = sorted((float(x), int(l)) for s in shards for l, x in zip(*s.search(q, k, nprobes=nprobes)))[:k]
                    if world == 1:
                        assert [c[1] for c in cand] == wl.tolist()
                    else:
                        assert (np.array([c[0] for c in cand], np.float32) <= wd * (1 + 1e-6)).all()
            finally:
                for s in shards:
                    s.close()


@pytest.mark.parametrize("metric,d", [("l2", 128), ("dot", 128), ("l2", 320), ("cosine", 320)])
def test_trainer_assignment_on_the_tensor_cores(gpu, oracle_mod, metric, d, monkeypatch):
    """Partition-level assignments of the trainer on the tensor cores (FlatGemm::nearest: bf16 tile maxima of the
    centroid table, exact fp32 distance to the rows of the four best tiles).  d=128 takes the TMEM-resident kernel (l2 as
    three extra bf16 columns), d=320 the K-streamed one (l2 as the bias epilogue; cosine trains on unit rows under l2).
    The exported assignment must be the nearest centroid; the index must search as well as the CUDA-core-trained one."""
    rng = np.random.default_rng(1234 + d)
    n, P, M, k = 60_000, 256, 16, 10
    centers = rng.standard_normal((400, d), dtype=np.float32) * 2
    X = (centers[rng.integers(0, 400, n)] + rng.standard_normal((n, d), dtype=np.float32)).astype(np.float32)
    Q = (centers[rng.integers(0, 400, 32)] + rng.standard_normal((32, d), dtype=np.float32)).astype(np.float32)
    Xn = X / np.linalg.norm(X, axis=1, keepdims=True) if metric == "cosine" else X

    def build(tc):
        monkeypatch.setenv("LANCE_B200_TRAIN_TC", "1" if tc else "0")
        h = gpu.LanceHandle(d, metric)
        h.add_batch(X)
        l0 = h.stats()["kernel_launches"]
        h.create_index(P, M)
        return h, h.stats()["kernel_launches"] - l0

    def recall(h):
        L, _, C = h.search_batch(Q, k, nprobes=16, refine_factor=3)
        hits = 0
        for i in range(Q.shape[0]):
            wl, _, _ = oracle_mod.flat_search(X, Q[i], k, metric)
            hits += len(set(int(x) for x in L[i, :C[i]]) & set(int(x) for x in wl))
        return hits / (k * Q.shape[0])

    h_tc, launches_tc = build(True)
    h_cc, launches_cc = build(False)
    try:
        assert launches_tc != launches_cc, "the tensor-core assignment did not run"
        g = h_tc.get_ivfpq()
        off = g["part_offsets"]
        assert off[0] == 0 and off[-1] == n and sorted(g["row_ids"].tolist()) == list(range(n))
        assert (np.diff(off) > 0).sum() >= P // 2, "k-means collapsed"
        C64 = g["centroids"].astype(np.float64)
        exact = 0
        sample = rng.integers(0, n, 400)
        for pos in sample:
            row = int(g["row_ids"][pos])
            p = int(np.searchsorted(off, pos, side="right") - 1)
            x = Xn[row].astype(np.float64)
            sc = -(C64 @ x) if metric == "dot" else ((C64 - x) ** 2).sum(1)
            tol = 1e-5 * abs(sc.min()) + 1e-5
            exact += int(sc[p] <= sc.min() + tol)
            # never far from the nearest: a miss can only be a bf16 near-tie inside one tile of the centroid table
            assert sc[p] <= sc.min() + 0.02 * abs(sc.min()) + 1e-3, f"row {row}: partition {p} is not near the best"
        assert exact >= 0.99 * len(sample), f"only {exact} of {len(sample)} sampled rows sit in their nearest partition"
        r_tc, r_cc = recall(h_tc), recall(h_cc)
        print(f"trainer {metric} d={d}: {exact}/{len(sample)} sampled rows in their nearest partition, recall {r_tc:.3f} (tensor cores) / {r_cc:.3f} (CUDA cores)")
        assert r_tc >= r_cc - 0.1, f"recall {r_tc:.2f} (tensor cores) vs {r_cc:.2f} (CUDA cores)"
    finally:
        h_tc.close()
        h_cc.close()

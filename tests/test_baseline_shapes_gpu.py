"""Oracle parity at the FULL shapes of BASELINE.json's configurations — the shapes every published number is on.

  C2  flat cosine  1 M x 768,  k=100, B=4096  (tcgen05 K-streamed kernel)        + single queries (fp32 scan)
  C3  IVF_PQ       10 M x 768, P=4096, M=96, nprobes=32, k=10: GPU-trained index exported to the checker
  C4  flat dot     100 M x 128, k=10, B=1024  (tcgen05 TMEM-resident kernel, > 2^31 bytes of rows)
  C5  flat l2      10 M x 384, k=10, prefilter bitmaps of 80 % and 1 % selectivity, single queries and a batch

The tables are bench.py's own synthetic tables (bench.table_chunks: same seeds, generated on the device), so what is
checked here is what `bench.py --workload cN` times.  The table never sits on the host as a whole: each generated
chunk is appended to the handle, copied into one pinned buffer and fed to the oracle's chunked fp64 accumulator
(oracle.FlatAccumulator = oracle_flat_search over row chunks, OpenMP), which holds the truth top-k of the sampled
queries.  Bar (tests/parity.py): ids identical except where truth distances tie within 1e-5
relative, distances within 1e-4 relative; IVF_PQ: probed partitions and ADC candidates identical, ADC within 1e-4,
post-refine exact distances.  Everything goes through the C ABI (ctypes -> liblance_b200.so)."""
import numpy as np
import pytest

from parity import assert_topk_parity

pytestmark = pytest.mark.gpu


def _torch():
    import torch

    return torch, torch.device("cuda", 0)


def _truth_from_handle(oracle_mod, h, q, metric):
    """fp64 truth distance of a label, from the row as the device store holds it (lance_detached_get_vector)."""
    return lambda lab: oracle_mod.distance_f64(metric, q, h.get_vector(int(lab)))


def _load_table(h, bench, dim, rows, accs, masks=None, dele=None):
    """Generate bench.py's table on the device, append it to the handle and stream it through the accumulators.
    accs: list of (FlatAccumulator, allow_bitmap_or_None)."""
    torch, dev = _torch()
    sp = torch.cuda.current_stream().cuda_stream
    stage = torch.empty((bench.GEN_CHUNK, dim), dtype=torch.float32).pin_memory()
    h.reserve(rows)
    for pos, blk in bench.table_chunks(dim, 0, rows, dev):
        m = blk.shape[0]
        h.add_batch_device(blk.data_ptr(), m, dim, sp)
        stage[:m].copy_(blk)
        torch.cuda.synchronize()
        for acc, allow in accs:
            acc.add(stage[:m].numpy(), row0=pos, allow=allow, deleted=dele)
    del stage
    assert h.count() == rows


def _check_lists(L, D, C, k, n_expected=None):
    """Size-independent properties of every returned list: k results, ascending distances, no duplicate label."""
    assert (C == (k if n_expected is None else n_expected)).all()
    assert (np.diff(D, axis=1) >= 0).all()
    for j in range(L.shape[0]):
        assert len(set(L[j].tolist())) == L.shape[1]


def test_c4_flat_dot_100m_x_128_batch_1024_vs_oracle(gpu, oracle_mod):
    import bench

    torch, dev = _torch()
    torch.cuda.empty_cache()
    wl = bench.WORKLOADS["c4"]
    n, d, k, B, metric = wl["rows"], wl["dim"], wl["k"], wl["batch"], wl["metric"]
    Qh = bench.bench_queries(B, d, dev).cpu().numpy()
    sample = [0, 1, 127, 128, 511, 777, 1000, 1023]
    acc = oracle_mod.FlatAccumulator(Qh[sample], k, metric)
    with gpu.LanceHandle(d, metric) as h:
        _load_table(h, bench, d, n, [(acc, None)])
        want = acc.finish()
        st0 = h.stats()
        L, D, C = h.search_batch(Qh, k)  # lance_detached_search_batch, host buffers
        st1 = h.stats()
        assert st1["flat_tc_batches"] == st0["flat_tc_batches"] + 1, "the tcgen05 path did not run"
        assert st1["tc_fallback_queries"] == st0["tc_fallback_queries"]
        _check_lists(L, D, C, k)
        assert L.max() >= 2 ** 24  # rows far beyond 2^31 bytes into the table are reachable
        for (wl_, wd64), j in zip(want, sample):
            assert_topk_parity(L[j], D[j], wl_, wd64, _truth_from_handle(oracle_mod, h, Qh[j], metric), what=f"c4 q{j}")
        # the single-query call (fp32 streaming scan over the same 51.2 GB) agrees with the oracle too
        for (wl_, wd64), j in list(zip(want, sample))[:2]:
            l1, d1 = h.search(Qh[j], k)
            assert_topk_parity(l1, d1, wl_, wd64, _truth_from_handle(oracle_mod, h, Qh[j], metric), what=f"c4 scan q{j}")
    torch.cuda.empty_cache()


def test_c2_flat_cosine_1m_x_768_k100_batch_4096_vs_oracle(gpu, oracle_mod):
    import bench

    torch, dev = _torch()
    torch.cuda.empty_cache()
    wl = bench.WORKLOADS["c2"]
    n, d, k, B, metric = wl["rows"], wl["dim"], wl["k"], wl["batch"], wl["metric"]
    Qh = bench.bench_queries(B, d, dev).cpu().numpy()
    sample = [0, 5, 127, 128, 2047, 2048, 3000, 4095]
    acc = oracle_mod.FlatAccumulator(Qh[sample], k, metric)
    with gpu.LanceHandle(d, metric) as h:
        _load_table(h, bench, d, n, [(acc, None)])
        want = acc.finish()
        st0 = h.stats()
        L, D, C = h.search_batch(Qh, k)
        st1 = h.stats()
        assert st1["flat_tc_batches"] == st0["flat_tc_batches"] + 1, "the tcgen05 path did not run"
        assert st1["tc_fallback_queries"] == st0["tc_fallback_queries"]
        _check_lists(L, D, C, k)
        for (wl_, wd64), j in zip(want, sample):
            assert_topk_parity(L[j], D[j], wl_, wd64, _truth_from_handle(oracle_mod, h, Qh[j], metric), what=f"c2 q{j}")
        # B = 64 (one query tile, HBM-bound regime) and B = 1 (fp32 scan) return the same lists
        L64, D64, C64 = h.search_batch(Qh[:64], k)
        for (wl_, wd64), j in zip(want[:2], sample[:2]):
            assert_topk_parity(L64[j], D64[j], wl_, wd64, _truth_from_handle(oracle_mod, h, Qh[j], metric), what=f"c2 b64 q{j}")
            l1, d1 = h.search(Qh[j], k)
            assert_topk_parity(l1, d1, wl_, wd64, _truth_from_handle(oracle_mod, h, Qh[j], metric), what=f"c2 b1 q{j}")
    torch.cuda.empty_cache()


def test_c5_flat_l2_10m_x_384_prefilter_80pct_and_1pct_vs_oracle(gpu, oracle_mod):
    import bench

    torch, dev = _torch()
    torch.cuda.empty_cache()
    wl = bench.WORKLOADS["c5"]
    n, d, k, metric = wl["rows"], wl["dim"], wl["k"], wl["metric"]
    Qh = bench.bench_queries(64, d, dev).cpu().numpy()
    sample = list(range(8))
    masks = {sel: bench.bench_mask(n, sel, dev).cpu().numpy() for sel in (0.8, 0.01)}
    bitmaps = {sel: gpu.make_bitmap(m) for sel, m in masks.items()}
    accs = {sel: oracle_mod.FlatAccumulator(Qh[sample], k, metric) for sel in masks}
    with gpu.LanceHandle(d, metric) as h:
        _load_table(h, bench, d, n, [(accs[sel], bitmaps[sel]) for sel in masks])
        for sel in masks:
            want = accs[sel].finish()
            st0 = h.stats()
            for (wl_, wd64), j in zip(want, sample):  # the SQL rewrite issues single queries: fp32 scan with the bitmap
                l1, d1 = h.search(Qh[j], k, allow=bitmaps[sel])
                assert masks[sel][l1].all()
                assert_topk_parity(l1, d1, wl_, wd64, _truth_from_handle(oracle_mod, h, Qh[j], metric), what=f"c5 sel={sel} q{j}")
            st1 = h.stats()
            assert st1["flat_simt_batches"] == st0["flat_simt_batches"] + len(sample)
            # a batch of 64 filtered queries goes through the tcgen05 path (l2 folded into the contraction)
            L, D, C = h.search_batch(Qh, k, allow=bitmaps[sel])
            st2 = h.stats()
            assert st2["flat_tc_batches"] == st1["flat_tc_batches"] + 1 and st2["tc_fallback_queries"] == st1["tc_fallback_queries"]
            _check_lists(L, D, C, k)
            assert masks[sel][L.reshape(-1)].all()
            for (wl_, wd64), j in zip(want, sample):
                assert_topk_parity(L[j], D[j], wl_, wd64, _truth_from_handle(oracle_mod, h, Qh[j], metric), what=f"c5 batch sel={sel} q{j}")
    torch.cuda.empty_cache()


def _adc_of(ix, q, lab):
    """ADC distance of label `lab` under the checker's own LUT (for candidates the checker's list cut)."""
    pos = int(np.nonzero(ix.row_ids == lab)[0][0])
    p = int(np.searchsorted(ix.part_offsets, pos, side="right") - 1)
    lut = ix.lut(q, p)
    s = float(lut[np.arange(ix.M), ix.codes[pos]].astype(np.float64).sum())
    return {0: s, 1: 0.5 * s, 2: 1.0 + s}[ix.metric]


def _centroid_dist64(ix, q):
    c = ix.centroids.astype(np.float64)
    qq = q.astype(np.float64)
    if ix.metric == 1:
        qq = qq / np.linalg.norm(qq)
    if ix.metric == 2:
        return 1.0 - c @ qq
    return ((c - qq) ** 2).sum(1)


def test_c3_ivfpq_10m_x_768_p4096_m96_nprobes32_vs_oracle(gpu, oracle_mod):
    import bench

    torch, dev = _torch()
    torch.cuda.empty_cache()
    wl = bench.WORKLOADS["c3"]
    n, d, k, metric, P, M, nprobes = wl["rows"], wl["dim"], wl["k"], wl["metric"], wl["P"], wl["M"], wl["nprobes"]
    B = 4096
    Qh = bench.bench_queries(B, d, dev).cpu().numpy()
    with gpu.LanceHandle(d, metric) as h:
        _load_table(h, bench, d, n, [])
        h.create_index(P, M)  # lance_detached_create_index: GPU k-means + PQ training + encode
        g = h.get_ivfpq()     # "the same trained index": exported to the checker
        assert (g["P"], g["M"], g["n_indexed"], g["dim"]) == (P, M, n, d)
        off = g["part_offsets"]
        assert off[0] == 0 and off[-1] == n and (np.diff(off) >= 0).all()
        ix = oracle_mod.IvfPqIndex(metric, d, P, M, g["centroids"], g["codebook"], off, g["row_ids"], g["codes"], n_indexed=n)

        # Queries whose fp64 probe boundary (nprobes-th vs next centroid) and ADC cut (k-th vs next candidate) are
        # clear of fp32 rounding (gap > 1e-4 relative; fp32 noise is ~1e-6): a tie there is a legal swap under the
        # parity bar but would change the partition / candidate SETS, which this test compares element by element.
        picked = []
        for j in range(64):
            cd = np.sort(_centroid_dist64(ix, Qh[j]))
            if (cd[nprobes] - cd[nprobes - 1]) <= 1e-4 * abs(cd[nprobes]):
                continue
            w = ix.search_parts(None, Qh[j], k + 1, ix.probe(Qh[j], nprobes)[0], 0)
            if len(w["cand_adc"]) > k and (w["cand_adc"][k] - w["cand_adc"][k - 1]) <= 1e-4 * abs(w["cand_adc"][k]):
                continue
            picked.append(j)
            if len(picked) == 8:
                break
        assert len(picked) == 8

        def expect_final(q, cand_labels):
            """a14: exact fp64 distance of the candidates, re-sorted by (distance, label), cut to k."""
            ent = sorted((oracle_mod.distance_f64(metric, q, h.get_vector(int(l))), int(l)) for l in cand_labels)[:k]
            return np.array([e[1] for e in ent], np.int64), np.array([e[0] for e in ent], np.float64)

        L, D, C = h.search_batch(Qh, k, nprobes=nprobes, refine_factor=1)  # B = 4096: tiled probe, one CTA per query
        assert h.stats()["ivfpq_batches"] >= 1
        _check_lists(L, D, C, k)
        for j in picked:
            q = Qh[j]
            got = h.ivfpq_search_debug(q, k, nprobes, 1)
            wp, _ = ix.probe(q, nprobes)
            assert list(got["parts"]) == list(wp) or sorted(got["parts"]) == sorted(wp), f"c3 q{j}: probed partitions"
            if list(got["parts"]) != list(wp):  # order swaps only between centroids tying within 1e-5 relative
                cd = _centroid_dist64(ix, q)
                for a, b in zip(got["parts"], wp):
                    assert abs(cd[a] - cd[b]) <= 1e-5 * abs(cd[b])
            want = ix.search_parts(None, q, k, wp, 0)  # pre-refine candidates: the k best ADC keys of the probed lists
            adc_known = dict(zip(want["cand_labels"].tolist(), want["cand_adc"].tolist()))
            adc_truth = lambda lab: adc_known[lab] if lab in adc_known else _adc_of(ix, q, lab)
            scale = float(np.abs(want["cand_adc"]).max())
            assert list(got["cand_labels"]) == list(want["cand_labels"]), f"c3 q{j}: candidate ids"
            assert_topk_parity(got["cand_labels"], got["cand_adc"], want["cand_labels"], want["cand_adc"], adc_truth,
                               scale=scale, what=f"c3 q{j} ADC candidates")
            fl, fd = expect_final(q, want["cand_labels"])
            truth = _truth_from_handle(oracle_mod, h, q, metric)
            assert_topk_parity(got["labels"], got["dist"], fl, fd, truth, what=f"c3 q{j} post-refine")
            # the reference-facing single call and the query's slot in the B = 4096 batch return that same list
            l1, d1 = h.search(q, k, nprobes=nprobes, refine_factor=1)
            assert list(l1) == list(got["labels"])
            np.testing.assert_array_equal(d1, got["dist"])
            assert_topk_parity(L[j], D[j], fl, fd, truth, what=f"c3 q{j} in the batch of {B}")
    del g, ix
    torch.cuda.empty_cache()

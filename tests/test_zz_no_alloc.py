"""lance_b200_flat_search_noalloc (include/lance_b200.h): the allocation-free flat search, through the C ABI.

No counterpart in the reference.  Checked against the existing entry
point lance_b200_search_device on the same device-resident table, twice in a row (the static scratch keeps the previous
call's contents), with the tolerances of tests/parity.py.  Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest

from parity import assert_topk_parity

pytestmark = pytest.mark.gpu

METRIC_ID = {"l2": 0, "cosine": 1, "dot": 2}


def _existing(gpu, torch, h, q_t, k, allow_t=None, bits=0):
    dev = q_t.device
    L = torch.empty(k, dtype=torch.int64, device=dev)
    D = torch.empty(k, dtype=torch.float32, device=dev)
    C = torch.empty(1, dtype=torch.int32, device=dev)
    h.search_device(q_t.data_ptr(), 1, k, L.data_ptr(), D.data_ptr(), C.data_ptr(),
                    d_allow=allow_t.data_ptr() if allow_t is not None else 0, allow_bits=bits)
    torch.cuda.synchronize()
    c = int(C.item())
    return L[:c].cpu().numpy(), D[:c].cpu().numpy()


def _noalloc(gpu, torch, X_t, metric, q_t, k, allow_t=None, bits=0, label_base=0):
    dev = q_t.device
    L = torch.full((k,), -7, dtype=torch.int64, device=dev)
    D = torch.full((k,), -7.0, dtype=torch.float32, device=dev)
    C = torch.full((1,), -7, dtype=torch.int32, device=dev)
    gpu.flat_search_noalloc(0, X_t.data_ptr(), X_t.shape[0], X_t.shape[1], METRIC_ID[metric], q_t.data_ptr(), k,
                            L.data_ptr(), D.data_ptr(), C.data_ptr(), label_base=label_base,
                            d_allow=allow_t.data_ptr() if allow_t is not None else 0, allow_bits=bits,
                            stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    c = int(C.item())
    assert 0 <= c <= k
    Lc, Dc = L.cpu().numpy(), D.cpu().numpy()
    assert (Lc[c:] == -1).all() and np.isinf(Dc[c:]).all()
    return Lc[:c], Dc[:c]


@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_noalloc_matches_existing_entry_point_twice_in_a_row(gpu, oracle_mod, metric):
    import torch

    torch.cuda.set_device(0)
    g = torch.Generator(device="cuda:0")
    g.manual_seed(77)
    n, d = 50000, 96
    X = torch.randn((n, d), generator=g, device="cuda:0", dtype=torch.float32)
    Xh = X.cpu().numpy()
    with gpu.LanceHandle(d, metric) as h:
        h.add_batch_device(X.data_ptr(), n, d)
        # two calls in a row with DIFFERENT queries and k: the second finds the first's keys in the static scratch
        for call, k in enumerate((10, 100, 7)):
            q = torch.randn(d, generator=g, device="cuda:0", dtype=torch.float32)
            wl, wd = _existing(gpu, torch, h, q, k)
            gl, gd = _noalloc(gpu, torch, X, metric, q, k)
            qh = q.cpu().numpy()
            assert_topk_parity(gl, gd, wl, wd, lambda lab: oracle_mod.distance_f64(metric, qh, Xh[lab]),
                               what=f"{metric} call {call}")


def test_noalloc_prefilter_small_tables_and_label_base(gpu, oracle_mod):
    import torch

    torch.cuda.set_device(0)
    g = torch.Generator(device="cuda:0")
    g.manual_seed(78)
    d = 33  # not a multiple of 4: the scalar load path
    for n, k in [(1, 5), (31, 40), (257, 128), (5000, 16)]:
        X = torch.randn((n, d), generator=g, device="cuda:0", dtype=torch.float32)
        q = torch.randn(d, generator=g, device="cuda:0", dtype=torch.float32)
        mask = (torch.rand(n, generator=g, device="cuda:0") < 0.5).cpu().numpy()
        bm = gpu.make_bitmap(mask)
        bm_t = torch.from_numpy(bm.view(np.int32)).to("cuda:0")
        with gpu.LanceHandle(d, "l2") as h:
            h.set_label_base(1000)
            h.add_batch_device(X.data_ptr(), n, d)
            Xh, qh = X.cpu().numpy(), q.cpu().numpy()
            truth = lambda lab: oracle_mod.distance_f64("l2", qh, Xh[lab - 1000])
            for allow_t, bits in [(None, 0), (bm_t, bm.shape[0] * 32)]:
                wl, wd = _existing(gpu, torch, h, q, k, allow_t, bits)
                gl, gd = _noalloc(gpu, torch, X, "l2", q, k, allow_t, bits, label_base=1000)
                assert len(gl) == min(k, n if allow_t is None else int(mask.sum()))
                assert_topk_parity(gl, gd, wl, wd, truth, what=f"n={n} k={k} filtered={allow_t is not None}")


def test_noalloc_rejects_sizes_beyond_the_static_scratch(gpu):
    import torch

    X = torch.zeros((8, 4), device="cuda:0")
    q = torch.zeros(4, device="cuda:0")
    L = torch.empty(200, dtype=torch.int64, device="cuda:0")
    D = torch.empty(200, dtype=torch.float32, device="cuda:0")
    C = torch.empty(1, dtype=torch.int32, device="cuda:0")
    for k in (0, 129, 4096):
        with pytest.raises(gpu.LanceError, match="exceeds the static scratch"):
            gpu.flat_search_noalloc(0, X.data_ptr(), 8, 4, 0, q.data_ptr(), k, L.data_ptr(), D.data_ptr(), C.data_ptr())
    with pytest.raises(gpu.LanceError, match=r"rows must be in \[0, 2\^32\)"):
        gpu.flat_search_noalloc(0, X.data_ptr(), 1 << 32, 4, 0, q.data_ptr(), 4, L.data_ptr(), D.data_ptr(), C.data_ptr())
    with pytest.raises(gpu.LanceError, match="metric must be"):
        gpu.flat_search_noalloc(0, X.data_ptr(), 8, 4, 3, q.data_ptr(), 4, L.data_ptr(), D.data_ptr(), C.data_ptr())
    with pytest.raises(gpu.LanceError, match="null buffer"):
        gpu.flat_search_noalloc(0, X.data_ptr(), 8, 4, 0, 0, 4, L.data_ptr(), D.data_ptr(), C.data_ptr())


def test_noalloc_call_allocates_nothing(gpu):
    """Device memory in use (driver view: cudaMemGetInfo) does not move across warmed-up calls."""
    import torch

    g = torch.Generator(device="cuda:0")
    g.manual_seed(79)
    X = torch.randn((20000, 64), generator=g, device="cuda:0")
    q = torch.randn(64, generator=g, device="cuda:0")
    L = torch.empty(16, dtype=torch.int64, device="cuda:0")
    D = torch.empty(16, dtype=torch.float32, device="cuda:0")
    C = torch.empty(1, dtype=torch.int32, device="cuda:0")
    call = lambda: gpu.flat_search_noalloc(0, X.data_ptr(), 20000, 64, 0, q.data_ptr(), 16, L.data_ptr(), D.data_ptr(),
                                           C.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
    call()  # set-up (module load, symbol address)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(50):
        call()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 == free0 and int(C.item()) == 16

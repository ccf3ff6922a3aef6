"""World-size-2 (and 3) gloo tests of the multi-GPU host logic (duckdb-lancedb_b200/sharding.py): contiguous
label-range shards, replicated queries, one all_gather of the per-shard top-k lists, final select.  The local
search and the final select are the oracle's here (CPU); on the GPU box bench.py plugs the CUDA search and
lance_b200_merge_topk_device into the same functions."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, d, nq, k, metric, outdir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import importlib

    import oracle
    from __graft_entry__ import load_package

    load_package()
    sharding = importlib.import_module("duckdb_lancedb_b200.sharding")
    rng = np.random.default_rng(1234)
    X = rng.standard_normal((n, d), dtype=np.float32)
    X[n // 2] = X[3]  # an exact duplicate in another shard: the tie must resolve to the smaller label
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    Q[0] = X[3]
    deleted = np.zeros(n, bool)
    deleted[rng.choice(n, n // 50, replace=False)] = True
    deleted[[3, n // 2]] = False
    lo, hi = sharding.shard_range(n, rank, world)
    L = np.full((nq, k), -1, np.int64)
    D = np.full((nq, k), np.inf, np.float32)
    C = np.zeros(nq, np.int32)
    for j in range(nq):
        wl, wd32, wd64 = oracle.flat_search(X[lo:hi], Q[j], k, metric, None, oracle.make_bitmap(deleted[lo:hi]), label_base=lo)
        C[j] = len(wl)
        L[j, :len(wl)] = wl
        D[j, :len(wl)] = wd32

    def merge(GL, GD, GC, G, nq_, k_, ML, MD, MC):
        for j in range(nq_):
            ml, md = oracle.merge_topk(GL[:, j].numpy(), GD[:, j].numpy(), GC[:, j].numpy(), k_)
            MC[j] = len(ml)
            ML[j, :len(ml)] = torch.from_numpy(ml)
            MD[j, :len(md)] = torch.from_numpy(md)

    sg = sharding.ShardGather(world, nq, k, torch.device("cpu"))
    ML, MD, MC = sg.gather_and_merge(torch.from_numpy(L), torch.from_numpy(D), torch.from_numpy(C), merge)
    # every rank must hold the same global answer
    ref = [ML.clone(), MD.clone(), MC.clone()]
    for t in ref:
        dist.broadcast(t, src=0)
    assert torch.equal(ref[0], ML) and torch.equal(ref[2], MC)
    if rank == 0:
        np.savez(os.path.join(outdir, "merged.npz"), L=ML.numpy(), D=MD.numpy(), C=MC.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,metric", [(2, "l2"), (2, "dot"), (3, "cosine")])
def test_sharded_search_equals_single_shard(oracle_mod, tmp_path, world, metric):
    n, d, nq, k = 3001, 16, 6, 10
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n, d, nq, k, metric, str(tmp_path)), nprocs=world, join=True)
    got = np.load(tmp_path / "merged.npz")
    rng = np.random.default_rng(1234)
    X = rng.standard_normal((n, d), dtype=np.float32)
    X[n // 2] = X[3]
    Q = rng.standard_normal((nq, d), dtype=np.float32)
    Q[0] = X[3]
    deleted = np.zeros(n, bool)
    deleted[rng.choice(n, n // 50, replace=False)] = True
    deleted[[3, n // 2]] = False
    for j in range(nq):
        wl, wd32, wd64 = oracle_mod.flat_search(X, Q[j], k, metric, None, oracle_mod.make_bitmap(deleted))
        assert got["C"][j] == len(wl)
        assert got["L"][j, :len(wl)].tolist() == wl.tolist(), f"q{j}"
        np.testing.assert_allclose(got["D"][j, :len(wl)], wd64, rtol=1e-4, atol=1e-6)
    if metric == "l2":
        assert got["L"][0, 0] == 3 and got["L"][0, 1] == n // 2  # duplicate rows: smaller label first


def test_shard_ranges_partition_the_table(pkg):
    import importlib

    sharding = importlib.import_module("duckdb_lancedb_b200.sharding")
    for rows in (0, 1, 7, 100_000_000, 12345):
        for world in (1, 2, 3, 4, 8):
            r = [sharding.shard_range(rows, g, world) for g in range(world)]
            assert r[0][0] == 0 and r[-1][1] == rows
            assert all(r[g][1] == r[g + 1][0] for g in range(world - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1

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


class _StubHandle:
    """Stands in for LanceHandle on CPU: the asynchronous search of rank `bad_rank` reports an overflowed query and
    leaves a wrong list for it until search_finish repairs it (what FlatGemm::finish does on the device)."""

    def __init__(self, rank, bad_rank, L, D, C, sg):
        self.rank, self.bad, self.Lt, self.Dt, self.Ct, self.sg = rank, bad_rank, L, D, C, sg
        self.finished = 0

    def search_device_async(self, d_queries, nq, k, d_labels, d_dist, d_counts, stream=0, d_overflow=0, **kw):
        self.sg.L.copy_(self.Lt)
        self.sg.D.copy_(self.Dt)
        self.sg.C.copy_(self.Ct)
        self.sg.O.zero_()
        if self.rank == self.bad:
            self.sg.C[1] = 0  # query 1 not final yet
            self.sg.O[0] = 1

    def search_finish(self, stream=0):
        self.finished += 1
        if self.rank == self.bad:
            self.sg.C[1] = self.Ct[1]
            return 1
        return 0


def _settle_worker(rank, world, port, outdir):
    import importlib
    import torch
    import torch.distributed as dist

    sys.path.insert(0, ROOT)
    from __graft_entry__ import load_package

    load_package()
    sharding = importlib.import_module("duckdb_lancedb_b200.sharding")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    nq, k = 3, 4
    # shard r holds distances r + 0.1*i, labels 100*r + i
    L = torch.tensor([[100 * rank + i for i in range(k)]] * nq, dtype=torch.int64)
    D = torch.tensor([[rank + 0.1 * i for i in range(k)]] * nq, dtype=torch.float32)
    C = torch.full((nq,), k, dtype=torch.int32)

    def merge(GL, GD, GC, G, nq_, k_, ML, MD, MC):
        for j in range(nq_):
            ents = sorted((float(GD[g, j, i]), int(GL[g, j, i])) for g in range(G) for i in range(int(GC[g, j])))[:k_]
            MC[j] = len(ents)
            for i, (dd, ll) in enumerate(ents):
                ML[j, i], MD[j, i] = ll, dd

    ss = sharding.ShardedSearch(None, None, world, nq, k, torch.device("cpu"), 0, merge=merge)
    ss.h = _StubHandle(rank, 0, L, D, C, ss.g)
    ML, MD, MC = ss.step(0, 0)
    stale = ML.clone()
    n = ss.settle()
    assert n == 1 and ss.h.finished == 1  # every rank settles, also the ones without an overflow of their own
    assert stale[1].tolist() == [100, 101, 102, 103]  # before: shard 0's list for query 1 was missing
    assert ML[1].tolist() == [0, 1, 2, 3] and ML[0].tolist() == [0, 1, 2, 3]
    assert ss.settle() == 0
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_step_settles_overflowed_shard(pkg, tmp_path):
    port = _free_port()
    mp.spawn(_settle_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)


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

#!/usr/bin/env python
"""bench.py — kNN queries/sec of the B200 search hot path (BASELINE.json metric), one JSON line.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (oracle port), rank 0 only
  torchrun --nproc-per-node N bench.py --gpus N ...          # N > 1: one process per GPU, NCCL gather + device merge

Headline workload (default `c4`, the configuration BASELINE.json's metric "queries/sec (k=10) at 1/2/4/8 B200" is
quoted on, and it fits one GPU): flat dot-product search, 100M x 128 fp32 rows sharded by contiguous row ranges over
the N GPUs, batch of 1024 queries replicated to every GPU, k=10.  A "step" = one batch through the hot path: local
top-k on every shard, all_gather of the per-shard lists, final device-side select.  Strong scaling (total rows fixed).

At N=1 the default run then times BASELINE.json's other configurations as well (c1; c2 at batch 1 / 64 / 4096; c3 at
batch 1 / 4096; c5 with the 80 % and the 1 % prefilter) and reports them under `configs` in the same JSON line, each
with value, ms_per_step, e2e, BOTH roofline fractions (step level = algorithmic work / ms_per_step, kernel level =
algorithmic work / CUDA-event time of the dominant kernel) and a CPU figure.  `--workload cN` times one of them as the
headline instead (no `configs`), `--no-configs` skips them.

`value`   : queries/s with queries, rows and results resident in HBM (CUDA events, max over ranks).
`e2e`     : same metric through the reference-facing C-ABI call with HOST buffers: pinned host queries are
            copied H2D and the final (label, distance) lists D2H inside the timed region, every step.
`roofline`: the dominant kernel's algorithmic bytes (or flops) / its CUDA-event duration, vs MEASURED_PEAKS.json.
`cpu_baseline`: the oracle's OpenMP fp32 search (cache-blocked SGEMM-style kernel for batches) timed on this box's
            host cores, on the full table where that takes seconds, else on a bounded sample (said in `sample`).
"""
from __future__ import annotations

import argparse
import json
import os
import math
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: rows, dim, metric, batch, k, kind
    "c1": dict(rows=100_000, dim=128, metric="l2", batch=1, k=10, kind="flat",
               desc="lance_search flat L2 k=10, 100Kx128 fp32, single query"),
    "c2": dict(rows=1_000_000, dim=768, metric="cosine", batch=4096, k=100, kind="flat",
               desc="flat cosine k=100, 1Mx768 fp32, query batch 4096 (also 1/64 via --batch)"),
    "c3": dict(rows=10_000_000, dim=768, metric="l2", batch=1, k=10, kind="ivfpq", P=4096, M=96, nprobes=32,
               desc="IVF_PQ (4096 partitions, 96 sub-quantizers) 10Mx768, nprobes=32, k=10"),
    "c4": dict(rows=100_000_000, dim=128, metric="dot", batch=1024, k=10, kind="flat",
               desc="flat dot-product 100Mx128 fp32 sharded by rows across the GPUs, batch 1024, k=10"),
    "c5": dict(rows=10_000_000, dim=384, metric="l2", batch=1, k=10, kind="flat", selectivity=0.8,
               desc="ORDER BY array_distance LIMIT 10 with WHERE prefilter (80% selectivity bitmap), 10Mx384"),
}
# the configurations the default N=1 run reports under `configs`, grouped by table: (name, workload, overrides)
SUBCONFIGS = [
    ("c1", "c1", {}),
    ("c2b1", "c2", {"batch": 1}), ("c2b64", "c2", {"batch": 64}), ("c2b4096", "c2", {"batch": 4096}),
    ("c5", "c5", {"selectivity": 0.8}), ("c5s01", "c5", {"selectivity": 0.01}),
    ("c3b1", "c3", {"batch": 1}), ("c3b4096", "c3", {"batch": 4096}),
]
SEED = 0xC0FFEE
GEN_CHUNK = 1 << 22


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=d["hbm_gbs"], bf16_tflops=d["bf16_tflops"],
                    bf16_tflops_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]), source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, source="fallback")


def load_traffic(name):
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of the dominant kernel from the committed
    `ncu --set full` capture of this configuration (profiles/traffic.json, written from profiles/*_ncu_full.txt)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None, None
    ent = json.load(open(p)).get(name)
    if not ent:
        return None, None
    return ent.get("bytes"), ent.get("note")


# ------------------------------------------------------------------------------------------------ synthetic data
def _mixture_centres(dim, dev):
    """--data clustered: 1 000 unit-norm centres (unit-norm mixture of 1 000 Gaussians, sigma 0.05)."""
    import torch

    g = torch.Generator(device=dev)
    g.manual_seed(SEED + 3)
    c = torch.randn((1000, dim), generator=g, device=dev, dtype=torch.float32)
    return c / c.norm(dim=1, keepdim=True)


def _clustered(n, dim, g, dev):
    import torch

    cen = _mixture_centres(dim, dev)
    idx = torch.randint(0, cen.shape[0], (n,), generator=g, device=dev)
    x = cen[idx] + 0.05 * torch.randn((n, dim), generator=g, device=dev, dtype=torch.float32)
    return x / x.norm(dim=1, keepdim=True)


def table_chunks(dim, lo, hi, dev, data="gaussian"):
    """The synthetic table, rows [lo, hi): yields (first_row, tensor[rows, dim]) pieces generated on the device.
    Chunk c of 2^22 rows is seeded with SEED + 7919*c, so every sharding (and tests/test_baseline_shapes_gpu.py)
    sees the same table."""
    import torch

    g = torch.Generator(device=dev)
    pos = lo
    while pos < hi:
        cidx = pos // GEN_CHUNK
        g.manual_seed(SEED + 7919 * cidx)
        if data == "clustered":
            blk = _clustered(GEN_CHUNK, dim, g, dev)
        else:
            blk = torch.randn((GEN_CHUNK, dim), generator=g, device=dev, dtype=torch.float32)
        a, b = pos - cidx * GEN_CHUNK, min(hi, (cidx + 1) * GEN_CHUNK) - cidx * GEN_CHUNK
        yield pos, blk[a:b]
        pos += b - a
        del blk


def bench_queries(B, dim, dev, data="gaussian"):
    import torch

    g = torch.Generator(device=dev)
    g.manual_seed(SEED + 1)
    if data == "clustered":
        return _clustered(B, dim, g, dev)
    return torch.randn((B, dim), generator=g, device=dev, dtype=torch.float32)


def bench_mask(rows, selectivity, dev):
    """Prefilter of the C5 workload: bool[rows], True = the row may be returned."""
    import torch

    g = torch.Generator(device=dev)
    g.manual_seed(SEED + 2)
    return torch.rand(rows, generator=g, device=dev) < selectivity


def bench_scores(rows, dev):
    """The C5 table's extra column `score INT ~ U[0,100)`, drawn from the same stream as bench_mask so
    that `score < round(100 s)` selects exactly the rows bench_mask(rows, s) keeps (s a multiple of 0.01)."""
    import torch

    g = torch.Generator(device=dev)
    g.manual_seed(SEED + 2)
    return (torch.rand(rows, generator=g, device=dev).double() * 100.0).floor().to(torch.int64)


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device = device
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def summary(self, t0=0.0, t1=float("inf")):
        """Samples that arrived in [t0, t1] (a timed region) are the ones reported."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, ln in list(self.lines):
            if t < t0 or t > t1:
                continue
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}

    def stop(self):
        if not self.proc:
            return
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()


def shard_range(rows, rank, world):
    """Contiguous row ranges (duckdb-lancedb_b200/sharding.py)."""
    return rows * rank // world, rows * (rank + 1) // world


# ------------------------------------------------------------------------------------------------ CPU arm
def host_threads(oracle):
    """All the host threads the CPU arm can use.  torchrun exports OMP_NUM_THREADS=1 to its workers, which would
    make the reference arm single-threaded at N > 1; the affinity mask of the process is what really bounds it."""
    try:
        avail = len(os.sched_getaffinity(0))
    except AttributeError:
        avail = os.cpu_count() or 1
    return max(oracle.num_threads(), avail)


def cpu_flat_kernel(oracle, nq):
    """Batches: the cache-blocked SGEMM-style kernel (6 x 16 register tiles); a few queries: the
    row-streaming kernel (a query panel of 16 would waste 15 lanes)."""
    if nq >= 8:
        return oracle.flat_search_batch_blocked, "oracle_flat_search_batch_blocked (cache-blocked SGEMM-style fp32, OpenMP)"
    return oracle.flat_search_batch_f32, "oracle_flat_search_batch_f32 (row-streaming fp32, OpenMP)"


class CpuFlatSample:
    """A host table (the full one, or a bounded sample of it) and the timed CPU search over it."""

    def __init__(self, wl, X, Q, allow, full_rows, threads):
        import oracle

        self.oracle, self.wl, self.X, self.Q, self.allow, self.threads = oracle, wl, X, Q, allow, threads
        self.full_rows = full_rows
        self.fn, self.fn_name = cpu_flat_kernel(oracle, Q.shape[0])
        self.fn(X[:1000], Q[:1], wl["k"], wl["metric"], nthreads=threads)  # warm the thread pool

    def step(self):
        t0 = time.perf_counter()
        self.fn(self.X, self.Q, self.wl["k"], self.wl["metric"], allow=self.allow, nthreads=self.threads)
        return time.perf_counter() - t0

    @property
    def scale(self):
        """time of a full step / time of a sample step: one pass over X per batch, every query of the batch."""
        return (self.full_rows / self.X.shape[0]) * (self.wl["batch"] / self.Q.shape[0])

    def describe(self):
        r, q = self.X.shape[0], self.Q.shape[0]
        if r == self.full_rows and q == self.wl["batch"]:
            return f"{self.fn_name} on the FULL table ({r} rows x {q} queries per step), {self.threads} threads"
        return (f"{self.fn_name} on {r} of {self.full_rows} rows x {q} of {self.wl['batch']} queries per step, "
                f"{self.threads} threads; scaled x{self.scale:.4g} (time is linear in rows x queries)")


def make_cpu_flat_sample(wl, threads, budget_flop=6.0e11, X_full=None, Q_full=None, allow_full=None, seed=SEED):
    """Sample shape: all queries of the batch (so the query-panel reuse of the blocked kernel is the real one) and as
    many rows as `budget_flop` allows, at least 64 K and at most 2 GiB of fp32 (4 GiB when the device table was fetched;
    larger than any host cache).  When the
    caller hands in the device table (X_full), the sample is its leading rows; else it is generated on the host."""
    import oracle

    B, d = wl["batch"], wl["dim"]
    q = min(B, 4096)
    rows = int(max(65_536, budget_flop / (2.0 * q * d)))
    rows = min(rows, wl["rows"], ((4 << 30) if X_full is not None else (2 << 30)) // (4 * d))
    if X_full is not None and X_full.shape[0] >= rows:
        X = X_full[:rows]
        Q = Q_full[:q]
        allow = None if allow_full is None else allow_full[: (rows + 31) // 32].copy()
    else:
        rng = np.random.default_rng(seed)
        X = rng.standard_normal((rows, d), dtype=np.float32)
        Q = rng.standard_normal((q, d), dtype=np.float32)
        allow = oracle.make_bitmap(rng.random(rows) < wl["selectivity"]) if wl.get("selectivity") else None
    if allow is not None and rows % 32:
        allow[-1] &= np.uint32((1 << (rows % 32)) - 1)
    return CpuFlatSample(wl, X, Q, allow, wl["rows"], threads)


_IVF_CACHE = {}


def cpu_ivfpq_sample(wl, nq, threads, seed=SEED):
    """The oracle's IVF_PQ search (probe -> residual LUT per probed list -> ADC scan -> refine) on an index of the
    workload's full shape (P lists, N rows of M-byte codes).  Codebooks / codes are random: the arithmetic and
    the bytes scanned per query are those of a trained index, only recall is meaningless.  Queries run
    concurrently, one per host thread.  Returns (seconds, queries)."""
    import oracle
    from concurrent.futures import ThreadPoolExecutor

    key = (wl["rows"], wl["dim"], wl["P"], wl["M"])
    if key not in _IVF_CACHE:
        rng = np.random.default_rng(seed)
        n, d, P, M = wl["rows"], wl["dim"], wl["P"], wl["M"]
        nx = 100_000  # rows kept for the refine step; labels wrap into it
        X = rng.standard_normal((nx, d), dtype=np.float32)
        cen = rng.standard_normal((P, d), dtype=np.float32)
        cb = rng.standard_normal((M, 256, d // M), dtype=np.float32)
        off = (np.arange(P + 1, dtype=np.int64) * n) // P
        ids = np.arange(n, dtype=np.int64) % nx
        codes = rng.integers(0, 256, size=(n, M), dtype=np.uint8)
        ix = oracle.IvfPqIndex(wl["metric"], d, P, M, cen, cb, off, ids, codes, n_indexed=nx)
        _IVF_CACHE[key] = (ix, X)
    ix, X = _IVF_CACHE[key]
    rng = np.random.default_rng(seed + 1)
    Q = rng.standard_normal((nq, wl["dim"]), dtype=np.float32)
    ix.search(X, Q[0], wl["k"], wl["nprobes"], 1)
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda j: ix.search(X, Q[j], wl["k"], wl["nprobes"], 1), range(nq)))
    return time.perf_counter() - t0, nq


def cpu_baseline(wl, seconds_target=10.0, X_full=None, Q_full=None, allow_full=None):
    """{"value", "unit", "cores", "kind", "sample", ...}: the oracle port timed on this box's host cores."""
    import oracle

    oracle.build()
    threads = host_threads(oracle)
    if wl["kind"] == "ivfpq":
        dt, nq = cpu_ivfpq_sample(wl, 4 * threads, threads)
        nq = int(max(4 * threads, min(4096, nq * seconds_target / max(dt, 1e-3))))
        dt, nq = cpu_ivfpq_sample(wl, nq, threads)
        return {"value": nq / dt, "unit": "queries/s", "cores": threads, "kind": "port", "extrapolated": False,
                "sample": f"oracle_ivfpq_search (oracle/oracle.c) on a full-shape index (P={wl['P']}, M={wl['M']}, "
                          f"{wl['rows']} coded rows, random codebooks/codes), {nq} queries, one per host thread, {dt:.2f} s"}
    s = make_cpu_flat_sample(wl, threads, budget_flop=3.0e11, X_full=X_full, Q_full=Q_full, allow_full=allow_full)
    dt, reps = s.step(), 1
    while dt < 0.5 * seconds_target and reps < 4000 and dt / reps * (reps + 1) < seconds_target:
        dt += s.step()
        reps += 1
    per_step = dt / reps
    qps = wl["batch"] / (per_step * s.scale)
    return {"value": qps, "unit": "queries/s", "cores": threads, "kind": "port", "extrapolated": s.scale != 1.0,
            "scale_factor": s.scale, "sample_rows": int(s.X.shape[0]), "sample_queries": int(s.Q.shape[0]),
            "sample_seconds_per_step": per_step, "sample": s.describe() + f"; {reps} repetitions, {dt:.2f} s"}


def run_reference(args, wl):
    """`--impl reference`: the reference's CPU implementation of the path on the host cores.  LanceDB itself cannot
    be built here (no Rust toolchain, crates not vendored: DESIGN.md section 2), so this is the oracle port
    (`kind: "port"`), all host threads.  A step is a BOUNDED SAMPLE of the workload (all queries of the batch, a row
    range); `ms_per_step` is what was actually timed per step, `value` the full-workload rate it implies
    (`extrapolated`, `scale_factor`)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle

    oracle.build()
    threads = host_threads(oracle)
    common = {"impl": "reference", "metric": "kNN queries/sec", "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps,
              "warmup": args.warmup, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
              "data": "synthetic N(0,1), seeded", "config": workload_config(wl, args)}
    if wl["kind"] == "ivfpq":
        times = []
        nq = 8 * threads
        for i in range(args.warmup + args.steps):
            dt, _ = cpu_ivfpq_sample(wl, nq, threads)
            if i >= args.warmup:
                times.append(dt)
        dt = sum(times) / len(times)
        qps = nq / dt
        sample = (f"oracle IVF_PQ port on a full-shape index (P={wl['P']}, M={wl['M']}, {wl['rows']} coded rows, random "
                  f"codebooks/codes), {nq} queries per step, one per host thread (LanceDB itself cannot be built here)")
        line = dict(common, value=qps, ms_per_step=1e3 * dt, extrapolated=False, queries_per_step=nq,
                    cpu_baseline={"value": qps, "unit": "queries/s", "cores": threads, "kind": "port", "sample": sample},
                    e2e={"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
        print(json.dumps(line))
        return
    s = make_cpu_flat_sample(wl, threads, budget_flop=6.0e11)
    times = []
    for i in range(args.warmup + args.steps):
        dt = s.step()
        if i >= args.warmup:
            times.append(dt)
    dt = sum(times) / len(times)
    qps = wl["batch"] / (dt * s.scale)
    sample = s.describe() + " (LanceDB itself cannot be built here: no Rust toolchain, crates not vendored)"
    line = dict(common, value=qps, ms_per_step=1e3 * dt, extrapolated=s.scale != 1.0, scale_factor=s.scale,
                sample_rows=int(s.X.shape[0]), sample_queries=int(s.Q.shape[0]),
                full_step_ms_implied=1e3 * dt * s.scale,
                cpu_baseline={"value": qps, "unit": "queries/s", "cores": threads, "kind": "port", "sample": sample},
                e2e={"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
    print(json.dumps(line))


def workload_config(wl, args, name=None):
    cfg = {"workload": f"{name or args.workload}: {wl['desc']}", "rows": wl["rows"], "dim": wl["dim"], "metric": wl["metric"],
           "batch": wl["batch"], "k": wl["k"], "sharding": f"rows/{args.gpus}", "l2_hygiene": "inputs larger than L2"
           if wl["rows"] * wl["dim"] * 4 // max(args.gpus, 1) > 200e6 else "L2 flushed between steps"}
    if wl.get("selectivity"):
        cfg["selectivity"] = wl["selectivity"]
    if wl["kind"] == "ivfpq" and args.gpus > 1:
        cfg["sharding"] = f"inverted lists dealt whole to {args.gpus} GPUs, balanced by length; one global index"
    if getattr(args, "data", "gaussian") != "gaussian":
        cfg["data"] = args.data
    return cfg


# ------------------------------------------------------------------------------------------------ product arm
class Bench:
    """One process' view of a run: device, rank, the clock sampler."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist

        from __graft_entry__ import load_package

        self.torch, self.dist, self.args = torch, dist, args
        self.pkg = load_package()
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world == 1 and args.gpus > 1:
            raise SystemExit("launch N>1 with torchrun: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.saved_stdout = None
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            # NCCL writes its version banner to stdout; stdout must carry exactly one JSON line, so point fd 1 at
            # stderr until that line is printed
            sys.stdout.flush()
            self.saved_stdout = os.dup(1)
            os.dup2(2, 1)
            dist.init_process_group("nccl", device_id=self.dev)
        self.pkg.set_default_device(self.local)
        self.peaks = load_peaks()
        self.stream = torch.cuda.current_stream()
        self.sp = self.stream.cuda_stream
        self.sampler = ClockSampler(self.local)
        if self.rank == 0:
            self.sampler.start()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def emit(self, line):
        if self.saved_stdout is not None:
            sys.stdout.flush()
            os.dup2(self.saved_stdout, 1)
        print(json.dumps(line), flush=True)

    def close(self):
        self.sampler.stop()
        if self.world > 1:
            self.dist.destroy_process_group()


class Table:
    """A workload's table resident in HBM (this rank's shard), optionally with its IVF_PQ index."""

    def __init__(self, b: Bench, wl):
        torch = b.torch
        self.b, self.wl = b, wl
        rows, dim = wl["rows"], wl["dim"]
        self.sizes = None
        self.X_host = None  # host copy of a small table for the full-table CPU figure (fetched once per table)
        if wl["kind"] == "ivfpq" and b.world > 1:
            self._build_partition_shard(b, wl)
            return
        self.lo, self.hi = shard_range(rows, b.rank, b.world)
        self.n_local = self.hi - self.lo
        h = b.pkg.LanceHandle(dim, wl["metric"])
        h.set_label_base(self.lo)
        h.reserve(self.n_local)
        for pos, blk in table_chunks(dim, self.lo, self.hi, b.dev, b.args.data):
            h.add_batch_device(blk.data_ptr(), blk.shape[0], dim, b.sp)
        assert h.count() == self.n_local
        self.h = h
        if wl["kind"] == "ivfpq":
            t0 = time.perf_counter()
            h.create_index(wl["P"], wl["M"])
            self.index_build_s = time.perf_counter() - t0
            self.sizes = np.diff(h.ivfpq_part_offsets())
        torch.cuda.empty_cache()

    def _build_partition_shard(self, b, wl):
        """IVF_PQ at N > 1: shard BY PARTITION (duckdb-lancedb_b200/sharding.py).  Rank 0 trains the one
        global index on the whole table and broadcasts it in the exchange layout; every rank generates the same table
        (same seeds), keeps the rows of the inverted lists it owns under their global labels and installs the index
        with the other lists empty."""
        import importlib

        torch, dist = b.torch, b.dist
        sharding = importlib.import_module("duckdb_lancedb_b200.sharding")
        rows, dim = wl["rows"], wl["dim"]
        X = torch.empty((rows, dim), dtype=torch.float32, device=b.dev)
        for pos, blk in table_chunks(dim, 0, rows, b.dev, b.args.data):
            X[pos:pos + blk.shape[0]] = blk
        meta = [None]
        index = None
        t0 = time.perf_counter()
        if b.rank == 0:
            full = b.pkg.LanceHandle(dim, wl["metric"])
            full.reserve(rows)
            full.add_batch_device(X.data_ptr(), rows, dim, b.sp)
            full.create_index(wl["P"], wl["M"])
            index = full.get_ivfpq()
            full.close()
            torch.cuda.empty_cache()
            meta = [{k: index[k] for k in ("metric", "dim", "P", "M", "n_indexed")}]
        self.index_build_s = time.perf_counter() - t0
        dist.broadcast_object_list(meta, src=0)
        m = meta[0]
        shapes = {"centroids": ((m["P"], dim), np.float32), "codebook": ((m["M"], 256, dim // m["M"]), np.float32),
                  "part_offsets": ((m["P"] + 1,), np.int64), "row_ids": ((m["n_indexed"],), np.int64),
                  "codes": ((m["n_indexed"], m["M"]), np.uint8)}
        got = dict(m)
        for name, (shape, dt) in shapes.items():  # NCCL broadcast of the exchange arrays (codes: 0.96 GB at C3)
            t = torch.from_numpy(index[name]).to(b.dev) if b.rank == 0 else torch.empty(shape, dtype=getattr(torch, np.dtype(dt).name), device=b.dev)
            dist.broadcast(t, src=0)
            got[name] = t.cpu().numpy()
            del t
        h, owner = sharding.ivfpq_partition_shard(b.pkg, got, X, b.rank, b.world, wl["metric"], b.sp)
        del X
        torch.cuda.empty_cache()
        self.h = h
        self.lo, self.hi = 0, rows
        self.n_local = h.count()
        self.sizes = np.diff(h.ivfpq_part_offsets())
        self.partition_owner = owner

    def close(self):
        self.h.close()
        self.b.torch.cuda.empty_cache()


def measure(b: Bench, tab: Table, wl, name, steps, headline):
    """Times one configuration on a resident table.  Returns the JSON-line dict (rank 0) or None."""
    torch, dist, args, pkg, dev = b.torch, b.dist, b.args, b.pkg, b.dev
    world, rank, stream, sp, h = b.world, b.rank, b.stream, b.sp, tab.h
    rows, dim, metric, B, k = wl["rows"], wl["dim"], wl["metric"], wl["batch"], wl["k"]
    n_local, lo, hi = tab.n_local, tab.lo, tab.hi
    peaks = b.peaks
    Qd = bench_queries(B, dim, dev, args.data)
    Qh = Qd.cpu().pin_memory()

    d_allow, allow_bits, allow_host = 0, 0, None
    if wl.get("selectivity"):
        mask = bench_mask(rows, wl["selectivity"], dev)[lo:hi]
        allow_host = pkg.make_bitmap(mask.cpu().numpy())
        allow_t = torch.from_numpy(allow_host.view(np.int32)).to(dev)
        d_allow, allow_bits = allow_t.data_ptr(), allow_host.shape[0] * 32
    # N=1: the reference-facing call takes the pushed-down WHERE as a STRING (src/lance_index.cpp:452-453); the index's
    # extra column lives in HBM and the predicate is evaluated (once, then cached) on the device
    predicate = None
    if allow_host is not None and world == 1:
        predicate = "score < %d" % int(round(100 * wl["selectivity"]))
        if not h.has_extra_columns():
            h.set_column("score", bench_scores(rows, dev)[lo:hi].cpu().numpy(), first_label=lo)
        bm, bits, _, _ = h.predicate_bitmap(predicate, words=allow_host.shape[0])
        if bits != n_local or not np.array_equal(bm, allow_host):
            raise RuntimeError("device predicate bitmap differs from the bench mask")

    nprobes = wl.get("nprobes", 20)
    adc_bytes = ivf_query_bytes = None
    if wl["kind"] == "ivfpq":
        probed_rows = 0
        for j in range(min(B, 16)):
            dbg = h.ivfpq_search_debug(Qh[j].numpy(), k, nprobes, 1)
            probed_rows += int(tab.sizes[dbg["parts"]].sum())
        probed_rows /= min(B, 16)
        # per query: codes (M B/row) + row term (4) + row id (4) of the probed lists + the query's lookup table
        adc_bytes = probed_rows * (wl["M"] + 8) + wl["M"] * 256 * 4
        ivf_query_bytes = adc_bytes + wl["P"] * dim * 4 + wl["M"] * 256 * (dim // wl["M"]) * 4 + k * dim * 4

    L = torch.empty((B, k), dtype=torch.int64, device=dev)
    D = torch.empty((B, k), dtype=torch.float32, device=dev)
    C = torch.empty(B, dtype=torch.int32, device=dev)
    ss = None
    if world > 1:
        import importlib

        sharding = importlib.import_module("duckdb_lancedb_b200.sharding")
        ss = sharding.ShardedSearch(pkg, h, world, B, k, dev, b.local)
    Lh = torch.empty((B, k), dtype=torch.int64).pin_memory()
    Dh = torch.empty((B, k), dtype=torch.float32).pin_memory()
    Ch = torch.empty(B, dtype=torch.int32).pin_memory()
    Qstage = torch.empty_like(Qd)

    # the two C calls of a step with their ctypes arguments built once: what is timed is the library, not the harness
    call_device = h.bind_search_device(Qd.data_ptr(), B, k, L.data_ptr(), D.data_ptr(), C.data_ptr(), nprobes=nprobes,
                                       refine_factor=1, d_allow=d_allow, allow_bits=allow_bits, stream=sp)
    call_host_bitmap = h.bind_search_batch(Qh.numpy(), k, (Lh.numpy(), Dh.numpy(), Ch.numpy()), nprobes=nprobes,
                                           refine_factor=1, allow=allow_host)
    call_host = call_host_bitmap if predicate is None else h.bind_search_batch(
        Qh.numpy(), k, (Lh.numpy(), Dh.numpy(), Ch.numpy()), nprobes=nprobes, refine_factor=1, predicate=predicate)

    def step_device(q_ptr):
        """local top-k -> (N>1) one NCCL all_gather of the packed lists -> final device select -> settle."""
        if world > 1:
            out = ss.step(q_ptr, sp, nprobes=nprobes, refine_factor=1, d_allow=d_allow, allow_bits=allow_bits)
            ss.settle()
            return out
        call_device()
        return L, D, C

    def step_e2e():
        if world == 1:
            # the reference-facing call: host buffers in, host buffers out (lance_detached_search_batch)
            call_host()
        else:
            Qstage.copy_(Qh, non_blocking=True)
            oL, oD, oC = step_device(Qstage.data_ptr())
            Lh.copy_(oL, non_blocking=True)
            Dh.copy_(oD, non_blocking=True)
            Ch.copy_(oC, non_blocking=True)
            stream.synchronize()

    shard_bytes = n_local * dim * 4 if wl["kind"] == "flat" else 1 << 40
    need_flush = shard_bytes <= 200e6  # the 126 MB L2 would otherwise serve the table on every step after the first
    # 512 MB: between steps (outside the timed intervals) half of it is written and the other half read — the table is
    # evicted AND no dirty line is left for the timed step to write back
    flush_buf = torch.empty(512 << 20, dtype=torch.uint8, device=dev) if need_flush else None

    wall_ms = [0.0]
    per_step_ms = []  # device time of every timed step, where steps are timed one by one (L2 flushed between them)

    def flush_l2():
        flush_buf[: 256 << 20].fill_(1)
        flush_buf[256 << 20:].view(torch.int32).sum()
        torch.cuda.synchronize()

    def timed(fn, nsteps, host_call=False):
        """K steps bracketed by barrier + synchronize; device time from CUDA events on the launching stream (max over
        ranks), host wall-clock of the steps themselves in wall_ms[0].  host_call: fn is a synchronous host-buffer call
        (its wall-clock IS the end-to-end time; it runs on the library's own stream)."""
        b.barrier()
        if need_flush and world == 1 and not host_call:
            # one event pair per step, L2 flushed before each: the loop runs in C (lance_b200_time_search_device), so a
            # 30 us step is not measured through the interpreter's call overhead
            t0 = time.perf_counter()
            per = h.time_search_device(Qd.data_ptr(), B, k, L.data_ptr(), D.data_ptr(), C.data_ptr(), nsteps, nprobes=nprobes,
                                       refine_factor=1, d_allow=d_allow, allow_bits=allow_bits, stream=sp,
                                       d_flush=flush_buf.data_ptr(), flush_bytes=flush_buf.numel())
            wall_ms[0] = (time.perf_counter() - t0) * 1e3
            b.barrier()
            ms = float(per.sum())
            per_step_ms[:] = [float(v) for v in per]
        elif need_flush:
            evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(nsteps)]
            wall = 0.0
            for a, e in evs:
                flush_l2()
                a.record(stream)
                t0 = time.perf_counter()
                fn()
                wall += time.perf_counter() - t0
                e.record(stream)
                e.synchronize()
            b.barrier()
            ms = sum(a.elapsed_time(e) for a, e in evs)
            per_step_ms[:] = [a.elapsed_time(e) for a, e in evs]
            wall_ms[0] = wall * 1e3
        else:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            t0 = time.perf_counter()
            for _ in range(nsteps):
                fn()
            wall_ms[0] = (time.perf_counter() - t0) * 1e3
            e1.record(stream)
            b.barrier()
            ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms

    # ---- warm-up, then the timed region (device-resident).  nvidia-smi has been streaming since the process
    # started; only samples from the region on are reported.
    warmup = max(args.warmup, 3)
    # the FIRST step after ingest is reported on its own: a batched flat search builds the bf16 shadow of the table
    # inside it, workspaces are allocated, kernel attributes set
    ef0, ef1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ef0.record(stream)
    step_device(Qd.data_ptr())
    ef1.record(stream)
    ef1.synchronize()
    first_step_ms = ef0.elapsed_time(ef1)
    for _ in range(warmup - 1):
        step_device(Qd.data_ptr())
    b.barrier()
    st0 = h.stats()
    h.profile_enable(True)
    h.profile_read()
    if ss:
        ss.merges = 0
    t_region = time.perf_counter()
    ms = timed(lambda: step_device(Qd.data_ptr()), steps)
    step_times = sorted(per_step_ms)
    prof = h.profile_read()
    h.profile_enable(False)
    st1 = h.stats()
    merges_value = ss.merges if ss else 0
    # ---- end-to-end (host buffers) timed region
    for _ in range(2):
        step_e2e()
    ms_e2e_dev = timed(step_e2e, steps, host_call=True)
    # N=1: the reference-facing call is synchronous and runs on the library's own stream, so the host wall-clock of
    # the steps is the end-to-end time (the events on torch's stream only bound it from below)
    ms_e2e = max(ms_e2e_dev, wall_ms[0]) if world == 1 else ms_e2e_dev
    ms_e2e_bitmap = None
    if predicate is not None:  # the same call with the bitmap evaluated on the host and uploaded per call
        for _ in range(2):
            call_host_bitmap()
        t_dev = timed(call_host_bitmap, steps, host_call=True)
        ms_e2e_bitmap = max(t_dev, wall_ms[0])
    # A timed region shorter than a few sampler periods (single queries, small shards) yields no clock sample:
    # keep the same device step running, untimed, until the region + extension spans ~0.5 s of load (headline) or
    # ~0.25 s (sub-configurations).  The count is derived from the max-reduced times, so every rank runs the same
    # number of steps (the step has a collective).
    ext_steps = 0
    span = 500.0 if headline else 250.0
    if ms + ms_e2e < span:
        ext_steps = int(min(20000, math.ceil((span - ms - ms_e2e) / max(ms / steps, 1e-3))))
        for _ in range(ext_steps):
            step_device(Qd.data_ptr())
        b.barrier()
    clocks = b.sampler.summary(t_region, time.perf_counter()) if rank == 0 else None
    if clocks is not None:
        clocks["window"] = ("timed region" if not ext_steps else
                            f"timed region + {ext_steps} untimed identical steps (region shorter than {span / 1e3:g} s)")

    launches = st1["kernel_launches"] - st0["kernel_launches"] + merges_value
    qps = B * steps / (ms / 1e3)
    qps_e2e = B * steps / (ms_e2e / 1e3)
    step_s = ms / steps / 1e3

    # ---- roofline of the dominant kernel (algorithmic bytes / flops per launch) and of the
    # whole step (the same algorithmic work / ms_per_step)
    roof = None
    if prof["flat_gemm_launches"] > 0:
        t = prof["flat_gemm_ms"] / prof["flat_gemm_launches"] / 1e3
        # the launches of a step together = the whole batch against the whole shard: 2*B*n*d flops (one pass,
        # norms / re-rank excluded)
        flops = 2.0 * B * n_local * dim
        launches_per_step = prof["flat_gemm_launches"] / steps
        achieved = flops / launches_per_step / t / 1e12
        kk = dim + (3 if metric == "l2" else 0)  # K of the contraction (l2 carries -|x|^2/2 as three bf16 terms)
        nkb = (((kk + 15) // 16 * 16) + 63) // 64
        variant = "flat_gemm_ts_kernel" if nkb <= 4 else ("flat_gemm_tsk_kernel" if nkb <= 12 else "flat_gemm_kernel")
        ridge = B / 2.0  # flop per byte of bf16 X read once
        peak = peaks["bf16_tflops_sustained"]
        bound = "tensor" if ridge >= peak * 1e12 / (peaks["hbm_gbs"] * 1e9) else "hbm"
        if bound == "tensor":
            roof = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                    "step_achieved": flops / step_s / 1e12, "step_frac": flops / step_s / 1e12 / peak,
                    "traffic": None, "kernel": variant + " (tcgen05 bf16, all segments of a batch)",
                    "peak_source": peaks["source"] + " bf16 sustained"}
        else:
            bytes_step = n_local * dim * 2.0
            a = bytes_step / launches_per_step / t / 1e9
            roof = {"bound": "hbm", "achieved": a, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": a / peaks["hbm_gbs"],
                    "step_achieved": bytes_step / step_s / 1e9, "step_frac": bytes_step / step_s / 1e9 / peaks["hbm_gbs"],
                    "traffic": None, "kernel": variant + " (tcgen05 bf16, HBM-bound batch: bf16 shadow read once)",
                    "peak_source": peaks["source"]}
    elif prof["adc_scan_launches"] > 0:
        t = prof["adc_scan_ms"] / prof["adc_scan_launches"] / 1e3
        launches_per_step = prof["adc_scan_launches"] / steps
        a = adc_bytes * B / launches_per_step / t / 1e9
        # whole search: per query the scan bytes + refine rows; the centroid table and the codebook are read once per
        # BATCH ("amortised over B if batched")
        shared = wl["P"] * dim * 4 + wl["M"] * 256 * (dim // wl["M"]) * 4
        whole = ((ivf_query_bytes - shared) * B + shared) / step_s / 1e9
        roof = {"bound": "hbm", "achieved": a, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": a / peaks["hbm_gbs"],
                "step_achieved": whole, "step_frac": whole / peaks["hbm_gbs"],
                "traffic": None, "kernel": ("ivfpq1_kernel (one query: probe + lookup table + ADC scan + refine in one cooperative launch; bytes = the scan's)"
                                           if B == 1 else "adc_scan_kernel (codes + row terms + ids of the probed lists + LUT)"),
                "peak_source": peaks["source"], "avg_launch_ms": t * 1e3, "bytes_per_query_scan": adc_bytes,
                "bytes_per_query_whole_search": ivf_query_bytes,
                "step_bytes_note": "step level counts the whole search: per query scan bytes + refine rows, per batch centroid table + codebook"}
    elif prof["flat_scan_launches"] > 0:
        t = prof["flat_scan_ms"] / prof["flat_scan_launches"] / 1e3
        launches_per_step = prof["flat_scan_launches"] / steps
        sel = wl.get("selectivity") or 1.0
        # X read once per pass of <= 4 queries; masked rows are skipped before any load (s*N*D*4 + N/8)
        bytes_pass = n_local * dim * 4.0 * sel + (n_local / 8.0 if wl.get("selectivity") else 0.0)
        a = bytes_pass / t / 1e9
        roof = {"bound": "hbm", "achieved": a, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": a / peaks["hbm_gbs"],
                "step_achieved": bytes_pass * launches_per_step / step_s / 1e9,
                "step_frac": bytes_pass * launches_per_step / step_s / 1e9 / peaks["hbm_gbs"],
                "traffic": None, "kernel": ("flat_scan1_kernel (fp32, one query: TMA-staged rows, register top-k)" if B == 1 and dim % 4 == 0
                                           else "flat_scan_kernel (fp32, <=4 queries per pass)"), "peak_source": peaks["source"],
                "launches_per_step": launches_per_step, "avg_launch_ms": t * 1e3, "selectivity": sel}

    if roof is not None:
        if roof["bound"] == "hbm" and roof.get("achieved") is not None:
            # SURVEY.md 8d asks for both denominators: the measured copy bandwidth (read + write streams; a read-only
            # scan can exceed it) and the nominal 8 TB/s
            roof["peak_nominal"] = 8000.0
            roof["frac_nominal"] = roof["achieved"] / 8000.0
        # the committed capture is of the stock shape of the configuration; an overridden shape has no measured traffic
        stock = not args.rows and not args.k and args.data == "gaussian"
        roof["traffic"], note = load_traffic(name) if stock else (None, None)
        if note:
            roof["traffic_note"] = note
    if rank != 0:
        return None
    line = {"metric": "kNN queries/sec", "value": qps, "unit": "queries/s", "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32" if st1["flat_tc_batches"] == st0["flat_tc_batches"]
            else "bf16 candidates + f32 re-rank", "data": "synthetic N(0,1), seeded" if args.data == "gaussian" else
            "synthetic unit-norm mixture of 1000 Gaussians (sigma 0.05), seeded", "config": workload_config(wl, args, name),
            "clocks": clocks,
            "e2e": {"value": qps_e2e, "unit": "queries/s", "h2d_bytes_per_step": int(Qh.numel() * 4 + (allow_host.nbytes if allow_host is not None and world == 1 and predicate is None else 0)),
                    "d2h_bytes_per_step": int(Lh.numel() * 8 + Dh.numel() * 4 + Ch.numel() * 4), "ms_per_step": ms_e2e / steps},
            "gpu_launches": int(launches), "roofline": roof, "first_step_ms": first_step_ms,
            **({"step_ms_quantiles": {"min": step_times[0], "p50": step_times[len(step_times) // 2],
                                      "p99": step_times[min(len(step_times) - 1, int(0.99 * len(step_times)))],
                                      "max": step_times[-1], "n": len(step_times)}} if step_times else {}),
            **({"prefilter": {"predicate": predicate, "evaluated": "on the device over the HBM-resident column, bitmap cached per (predicate, epoch)",
                              "e2e_host_bitmap_ms_per_step": ms_e2e_bitmap / steps,
                              "host_bitmap_h2d_bytes_per_step": int(allow_host.nbytes)}} if predicate is not None else {}),
            "paths": {"flat_simt_batches": st1["flat_simt_batches"] - st0["flat_simt_batches"],
                      "flat_tc_batches": st1["flat_tc_batches"] - st0["flat_tc_batches"],
                      "tc_fallback_queries": st1["tc_fallback_queries"] - st0["tc_fallback_queries"],
                      "ivfpq_batches": st1["ivfpq_batches"] - st0["ivfpq_batches"]},
            "profile_ms": {kk: vv for kk, vv in prof.items()}}
    if wl["kind"] == "ivfpq":
        line["index_build_s"] = tab.index_build_s
    if not args.no_cpu_baseline and world == 1:
        # tables of up to 4 GiB come to the host whole (the CPU figure is then measured on the FULL table when a step
        # fits the budget); larger ones are sampled (all queries of the batch, a row range) and scaled
        X_full = Q_full = None
        if wl["kind"] == "flat" and rows * dim * 4 <= (4 << 30) and args.data == "gaussian":
            if tab.X_host is None:
                _, tab.X_host = h.get_all_vectors()
            X_full, Q_full = tab.X_host, Qh.numpy()
        line["cpu_baseline"] = cpu_baseline(wl, 10.0 if headline else 4.0, X_full, Q_full, allow_host)
    else:
        line["cpu_baseline"] = None
    return line


def run_b200(args, wl):
    b = Bench(args)
    tab = Table(b, wl)
    name = args.workload
    if args.batch and args.batch != WORKLOADS[args.workload]["batch"] or args.workload in ("c2", "c3"):
        name = f"{args.workload}b{wl['batch']}"
    if args.workload == "c5" and wl.get("selectivity") == 0.01:
        name = "c5s01"
    line = measure(b, tab, wl, name, args.steps, True)
    tab.close()
    configs = None
    if b.world == 1 and args.workload == "c4" and not args.no_configs and not (args.batch or args.rows or args.k):
        # BASELINE.json's other configurations, each on its own table, same protocol, fewer steps
        configs, cur, cur_wl = {}, None, None
        for name, base, over in SUBCONFIGS:
            sub = dict(WORKLOADS[base])
            sub.update(over)
            if "selectivity" in over:
                sub["desc"] = sub["desc"].replace("80% selectivity", f"{100 * over['selectivity']:g}% selectivity")
            try:
                if cur_wl != base:
                    if cur is not None:
                        cur.close()
                    cur, cur_wl = Table(b, sub), base
                steps = min(args.steps, 10 if sub["batch"] >= 1024 else 20)
                ln = measure(b, cur, sub, name, steps, False)
                for drop in ("metric", "unit", "n_gpus", "higher_is_better", "scaling", "vs_baseline", "data"):
                    ln.pop(drop, None)
                configs[name] = ln
            except Exception as e:  # a sub-configuration must never take the headline line down with it
                configs[name] = {"error": f"{type(e).__name__}: {e}"}
        if cur is not None:
            cur.close()
    if b.rank == 0:
        if configs is not None:
            line["configs"] = configs
        b.emit(line)
    b.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=None, help="override the workload's query batch")
    ap.add_argument("--rows", type=int, default=None, help="override the workload's row count (experiments only)")
    ap.add_argument("--k", type=int, default=None)
    ap.add_argument("--selectivity", type=float, default=None, help="c5: fraction of rows the prefilter keeps (0.8 / 0.01)")
    ap.add_argument("--data", default="gaussian", choices=["gaussian", "clustered"],
                    help="clustered: unit-norm mixture of 1000 Gaussians, sigma 0.05 (robustness of the candidate bound)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="N=1 default run: skip BASELINE.json's other configurations")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.batch:
        wl["batch"] = args.batch
    if args.rows:
        wl["rows"] = args.rows
    if args.k:
        wl["k"] = args.k
    if args.selectivity is not None and wl.get("selectivity"):
        wl["selectivity"] = args.selectivity
        wl["desc"] = wl["desc"].replace("80% selectivity", f"{100 * args.selectivity:g}% selectivity")
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_b200(args, wl)


if __name__ == "__main__":
    main()

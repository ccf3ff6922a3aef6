#!/usr/bin/env python
"""bench.py — kNN queries/sec of the B200 flat-search hot path (BASELINE.json metric), one JSON line.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (oracle port), rank 0 only
  torchrun --nproc-per-node N bench.py --gpus N ...          # N > 1: one process per GPU, NCCL gather + device merge

Workload (default `c4`, the configuration BASELINE.json's metric "queries/sec (k=10) at 1/2/4/8 B200" is quoted on,
and it fits one GPU): flat dot-product search, 100M x 128 fp32 rows sharded by contiguous row ranges over the
N GPUs, batch of 1024 queries replicated to every GPU, k=10.  A "step" = one batch through the hot path:
local top-k on every shard, all_gather of the per-shard lists, final device-side select.  Strong scaling
(total rows fixed).  Other workloads (`--workload c1|c2|c3|c5`) are BASELINE.json's remaining configs.

`value`   : queries/s with queries, rows and results resident in HBM (CUDA events, max over ranks).
`e2e`     : same metric through the reference-facing C-ABI call with HOST buffers: pinned host queries are
            copied H2D and the final (label, distance) lists D2H inside the timed region, every step.
`roofline`: the dominant kernel's algorithmic bytes (or flops) / its CUDA-event duration, vs MEASURED_PEAKS.json.
`cpu_baseline`: the oracle's OpenMP fp32 batch search timed on this box's host cores on a bounded sample.
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
SEED = 0xC0FFEE


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=d["hbm_gbs"], bf16_tflops=d["bf16_tflops"],
                    bf16_tflops_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]), source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, source="fallback")


def load_traffic(workload, batch):
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of the dominant kernel from the committed
    `ncu --set full` capture of this workload (profiles/traffic.json, written from profiles/*_ncu_full.txt)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None, None
    ent = json.load(open(p)).get(f"{workload}_b{batch}")
    if not ent:
        return None, None
    return ent.get("bytes"), ent.get("note")


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

    def count_since(self, t0):
        return sum(1 for t, _ in self.lines if t >= t0)

    def stop(self, t0=0.0):
        """Samples that arrived at or after t0 (the start of the timed region) are the ones reported."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, ln in self.lines:
            if t < t0:
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


GEN_CHUNK = 1 << 22


def table_chunks(dim, lo, hi, dev):
    """The synthetic table, rows [lo, hi): yields (first_row, tensor[rows, dim]) pieces generated on the device.
    Chunk c of 2^22 rows is seeded with SEED + 7919*c, so every sharding (and tests/test_baseline_shapes_gpu.py)
    sees the same table."""
    import torch

    g = torch.Generator(device=dev)
    pos = lo
    while pos < hi:
        cidx = pos // GEN_CHUNK
        g.manual_seed(SEED + 7919 * cidx)
        blk = torch.randn((GEN_CHUNK, dim), generator=g, device=dev, dtype=torch.float32)
        a, b = pos - cidx * GEN_CHUNK, min(hi, (cidx + 1) * GEN_CHUNK) - cidx * GEN_CHUNK
        yield pos, blk[a:b]
        pos += b - a
        del blk


def bench_queries(B, dim, dev):
    import torch

    g = torch.Generator(device=dev)
    g.manual_seed(SEED + 1)
    return torch.randn((B, dim), generator=g, device=dev, dtype=torch.float32)


def bench_mask(rows, selectivity, dev):
    """Prefilter of the C5 workload: bool[rows], True = the row may be returned."""
    import torch

    g = torch.Generator(device=dev)
    g.manual_seed(SEED + 2)
    return torch.rand(rows, generator=g, device=dev) < selectivity


def shard_range(rows, rank, world):
    """Contiguous row ranges (duckdb-lancedb_b200/sharding.py)."""
    return rows * rank // world, rows * (rank + 1) // world


# ------------------------------------------------------------------------------------------------ reference arm
_CPU_SAMPLE = {}


def cpu_search_sample(wl, sample_rows, sample_q, threads, seed=SEED):
    """The oracle's OpenMP fp32 flat search on a bounded sample; returns (seconds, rows, queries).  The sample
    (generated once per shape) is outside the timed region."""
    import oracle

    key = (sample_rows, sample_q, wl["dim"], wl.get("selectivity"))
    if key not in _CPU_SAMPLE:
        _CPU_SAMPLE.clear()
        rng = np.random.default_rng(seed)
        X = rng.standard_normal((sample_rows, wl["dim"]), dtype=np.float32)
        Q = rng.standard_normal((sample_q, wl["dim"]), dtype=np.float32)
        allow = None
        if wl.get("selectivity"):
            allow = oracle.make_bitmap(rng.random(sample_rows) < wl["selectivity"])
        oracle.flat_search_batch_f32(X[:1000], Q[:1], wl["k"], wl["metric"], nthreads=threads)  # warm the pool
        _CPU_SAMPLE[key] = (X, Q, allow)
    X, Q, allow = _CPU_SAMPLE[key]
    t0 = time.perf_counter()
    oracle.flat_search_batch_f32(X, Q, wl["k"], wl["metric"], allow=allow, nthreads=threads)
    return time.perf_counter() - t0, sample_rows, sample_q


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


def host_threads(oracle):
    """All the host threads the CPU arm can use.  torchrun exports OMP_NUM_THREADS=1 to its workers, which would
    make the reference arm single-threaded at N > 1; the affinity mask of the process is what really bounds it."""
    try:
        avail = len(os.sched_getaffinity(0))
    except AttributeError:
        avail = os.cpu_count() or 1
    return max(oracle.num_threads(), avail)


def cpu_sample_shape(wl, budget_pairs):
    """Pick (rows, queries) of the sample so that rows*queries*dim ~ budget, keeping the batch shape when possible."""
    q = min(wl["batch"], 64)
    rows = int(min(wl["rows"], max(10_000, budget_pairs // (q * wl["dim"]))))
    # bound the host memory of the sample (<= 2 GiB of fp32 rows, larger than any host cache); the time budget is
    # filled by repeating the sample instead
    rows = min(rows, (2 << 30) // (4 * wl["dim"]))
    return rows, q


def cpu_baseline(wl, seconds_target=12.0):
    import oracle

    oracle.build()
    threads = host_threads(oracle)
    if wl["kind"] == "ivfpq":
        dt, nq = cpu_ivfpq_sample(wl, 4 * threads, threads)
        nq = int(max(4 * threads, min(4096, nq * seconds_target / max(dt, 1e-3))))
        dt, nq = cpu_ivfpq_sample(wl, nq, threads)
        return {"value": nq / dt, "unit": "queries/s", "cores": threads, "kind": "port",
                "sample": f"oracle_ivfpq_search (oracle/oracle.c) on a full-shape index (P={wl['P']}, M={wl['M']}, "
                          f"{wl['rows']} coded rows, random codebooks/codes), {nq} queries, one per host thread, {dt:.2f} s"}
    # calibrate: ~0.3 s probe, then scale to the target
    rows, q = cpu_sample_shape(wl, 2_000_000_000)
    dt, r, qq = cpu_search_sample(wl, min(rows, 200_000), q, threads)
    rate = r * qq * wl["dim"] / dt
    rows, q = cpu_sample_shape(wl, int(rate * seconds_target))
    dt, r, qq = cpu_search_sample(wl, rows, q, threads)
    reps = 1
    while dt < 0.5 * seconds_target and reps < 4000:  # small tables / single queries: repeat the sample
        d2, _, _ = cpu_search_sample(wl, rows, q, threads)
        dt += d2
        reps += 1
    qq *= reps
    # queries/s on the full table: time scales with rows (one pass over X per batch)
    qps = qq / (dt * (wl["rows"] / r))
    return {"value": qps, "unit": "queries/s", "cores": threads, "kind": "port",
            "sample": f"oracle_flat_search_batch_f32 (OpenMP fp32 port, oracle/oracle.c) on {r} of {wl['rows']} rows x "
                      f"{qq} queries, {dt:.2f} s, scaled by rows"}


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle

    oracle.build()
    threads = host_threads(oracle)
    if wl["kind"] == "ivfpq":
        times = []
        nq = 8 * threads
        for i in range(args.warmup + args.steps):
            dt, _ = cpu_ivfpq_sample(wl, nq, threads)
            if i >= args.warmup:
                times.append(dt)
        qps = nq / (sum(times) / len(times))
        sample = (f"oracle IVF_PQ port on a full-shape index (P={wl['P']}, M={wl['M']}, {wl['rows']} coded rows, random "
                  f"codebooks/codes), {nq} queries per step, one per host thread (LanceDB itself cannot be built here)")
        line = {"impl": "reference", "metric": "kNN queries/sec", "value": qps, "unit": "queries/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wl["batch"] / qps,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": workload_config(wl, args),
                "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": threads, "kind": "port", "sample": sample},
                "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return
    rows, q = cpu_sample_shape(wl, 1_500_000_000)
    times, reps_per_step = [], 1
    for i in range(args.warmup + args.steps):
        dt, reps = 0.0, 0
        while reps < reps_per_step or (dt < 1.0 and reps < 2000):  # a step = at least ~1 s of CPU work
            d1, r, qq = cpu_search_sample(wl, rows, q, threads)
            dt += d1
            reps += 1
        reps_per_step = reps
        if i >= args.warmup:
            times.append(dt / reps)
    dt = sum(times) / len(times)
    qps = q / (dt * (wl["rows"] / rows))
    sample = (f"oracle OpenMP fp32 port of the reference's CPU flat search, {rows} of {wl['rows']} rows x {q} queries per "
              f"step, scaled by rows (LanceDB itself cannot be built here: no Rust toolchain, crates not vendored)")
    line = {"impl": "reference", "metric": "kNN queries/sec", "value": qps, "unit": "queries/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wl["batch"] / qps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(wl, args),
            "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(wl, args):
    return {"workload": f"{args.workload}: {wl['desc']}", "rows": wl["rows"], "dim": wl["dim"], "metric": wl["metric"],
            "batch": wl["batch"], "k": wl["k"], "sharding": f"rows/{args.gpus}", "l2_hygiene": "inputs larger than L2"
            if wl["rows"] * wl["dim"] * 4 // max(args.gpus, 1) > 200e6 else "L2 flushed between steps"}


# ------------------------------------------------------------------------------------------------ product arm
def run_b200(args, wl):
    import torch
    import torch.distributed as dist

    from __graft_entry__ import load_package

    pkg = load_package()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N>1 with torchrun: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    saved_stdout = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its version banner to stdout; stdout must carry exactly one JSON line, so point fd 1 at
        # stderr until that line is printed
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    pkg.set_default_device(local)
    peaks = load_peaks()

    rows, dim, metric, B, k = wl["rows"], wl["dim"], wl["metric"], wl["batch"], wl["k"]
    lo, hi = shard_range(rows, rank, world)
    n_local = hi - lo
    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream

    # ---- build the shard in HBM (synthetic N(0,1), seeded per chunk so any sharding sees the same table)
    h = pkg.LanceHandle(dim, metric)
    h.set_label_base(lo)
    h.reserve(n_local)
    for pos, blk in table_chunks(dim, lo, hi, dev):
        h.add_batch_device(blk.data_ptr(), blk.shape[0], dim, sp)
    assert h.count() == n_local
    Qd = bench_queries(B, dim, dev)
    Qh = Qd.cpu().pin_memory()

    d_allow, allow_bits, allow_host = 0, 0, None
    if wl.get("selectivity"):
        mask = bench_mask(rows, wl["selectivity"], dev)[lo:hi]
        allow_host = pkg.make_bitmap(mask.cpu().numpy())
        allow_t = torch.from_numpy(allow_host.view(np.int32)).to(dev)
        d_allow, allow_bits = allow_t.data_ptr(), allow_host.shape[0] * 32

    nprobes = wl.get("nprobes", 20)
    adc_bytes = None
    if wl["kind"] == "ivfpq":
        h.create_index(wl["P"], wl["M"])
        offs = h.ivfpq_part_offsets()
        sizes = np.diff(offs)
        probed_rows = 0
        for j in range(min(B, 16)):
            dbg = h.ivfpq_search_debug(Qh[j].numpy(), k, nprobes, 1)
            probed_rows += int(sizes[dbg["parts"]].sum())
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
        ss = sharding.ShardedSearch(pkg, h, world, B, k, dev, local)
    Lh = torch.empty((B, k), dtype=torch.int64).pin_memory()
    Dh = torch.empty((B, k), dtype=torch.float32).pin_memory()
    Ch = torch.empty(B, dtype=torch.int32).pin_memory()
    Qstage = torch.empty_like(Qd)

    def step_device(q_ptr):
        """local top-k -> (N>1) one NCCL all_gather of the packed lists -> final device select -> settle."""
        if world > 1:
            out = ss.step(q_ptr, sp, nprobes=nprobes, refine_factor=1, d_allow=d_allow, allow_bits=allow_bits)
            ss.settle()
            return out
        h.search_device(q_ptr, B, k, L.data_ptr(), D.data_ptr(), C.data_ptr(), nprobes=nprobes, refine_factor=1,
                        d_allow=d_allow, allow_bits=allow_bits, stream=sp)
        return L, D, C

    def step_e2e():
        if world == 1:
            # the reference-facing call: host buffers in, host buffers out (lance_detached_search_batch)
            h.search_batch(Qh.numpy(), k, nprobes=nprobes, refine_factor=1, allow=allow_host,
                           out=(Lh.numpy(), Dh.numpy(), Ch.numpy()))
        else:
            Qstage.copy_(Qh, non_blocking=True)
            oL, oD, oC = step_device(Qstage.data_ptr())
            Lh.copy_(oL, non_blocking=True)
            Dh.copy_(oD, non_blocking=True)
            Ch.copy_(oC, non_blocking=True)
            stream.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    shard_bytes = n_local * dim * 4 if wl["kind"] == "flat" else 1 << 40
    need_flush = shard_bytes <= 200e6  # the 126 MB L2 would otherwise serve the table on every step after the first
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if need_flush else None
    flush_rd = torch.zeros(64 << 20, dtype=torch.float32, device=dev) if need_flush else None  # 256 MB, read-only pass

    wall_ms = [0.0]

    def timed(fn, steps):
        """K steps bracketed by barrier + synchronize; device time from CUDA events on the launching stream (max over
        ranks), host wall-clock of the same steps in wall_ms[0]."""
        barrier()
        if need_flush:
            # one event pair per step.  Between steps (outside the timed intervals): write a buffer larger than L2, read
            # another one — the table is evicted AND the dirty lines of the write are already back in DRAM — and wait,
            # so that a step that runs on the library's own stream cannot overlap the flush
            evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            wall = 0.0
            for a, b in evs:
                flush_buf.fill_(1)
                flush_rd.sum()
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                a.record(stream)
                fn()
                b.record(stream)
                b.synchronize()
                wall += time.perf_counter() - t0
            barrier()
            ms = sum(a.elapsed_time(b) for a, b in evs)
            wall_ms[0] = wall * 1e3
        else:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record(stream)
            for _ in range(steps):
                fn()
            e1.record(stream)
            barrier()
            wall_ms[0] = (time.perf_counter() - t0) * 1e3
            ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms

    # ---- warm-up, then the timed region (device-resident).  nvidia-smi is started before the warm-up so that it is
    # already streaming when the timed region begins; only samples from the region on are reported.
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        step_device(Qd.data_ptr())
    barrier()
    st0 = h.stats()
    h.profile_enable(True)
    h.profile_read()
    if ss:
        ss.merges = 0
    t_region = time.perf_counter()
    ms = timed(lambda: step_device(Qd.data_ptr()), args.steps)
    prof = h.profile_read()
    h.profile_enable(False)
    st1 = h.stats()
    merges_value = ss.merges if ss else 0
    # ---- end-to-end (host buffers) timed region
    for _ in range(2):
        step_e2e()
    ms_e2e_dev = timed(step_e2e, args.steps)
    # N=1: the reference-facing call is synchronous and runs on the library's own stream, so the host wall-clock of
    # the steps is the end-to-end time (the events on torch's stream only bound it from below)
    ms_e2e = max(ms_e2e_dev, wall_ms[0]) if world == 1 else ms_e2e_dev
    # A timed region shorter than a few sampler periods (single queries, small shards) yields no clock sample:
    # keep the same device step running, untimed, until the region + extension spans ~0.5 s of load.  The count is
    # derived from the max-reduced times, so every rank runs the same number of steps (the step has a collective).
    ext_steps = 0
    if ms + ms_e2e < 500.0:
        ext_steps = int(min(20000, math.ceil((500.0 - ms - ms_e2e) / max(ms / args.steps, 1e-3))))
        for _ in range(ext_steps):
            step_device(Qd.data_ptr())
        barrier()
    clocks = sampler.stop(t_region) if rank == 0 else None
    if clocks is not None:
        clocks["window"] = ("timed region" if not ext_steps else
                            f"timed region + {ext_steps} untimed identical steps (region shorter than 0.5 s = five sampler periods)")

    launches = st1["kernel_launches"] - st0["kernel_launches"] + merges_value
    qps = B * args.steps / (ms / 1e3)
    qps_e2e = B * args.steps / (ms_e2e / 1e3)

    # ---- roofline of the dominant kernel (algorithmic bytes / flops per launch)
    roof = None
    if prof["flat_gemm_launches"] > 0:
        t = prof["flat_gemm_ms"] / prof["flat_gemm_launches"] / 1e3
        # one launch = the whole batch against the whole shard: 2*B*n*d flops (one pass, norms/re-rank excluded)
        flops = 2.0 * B * n_local * dim * (prof["flat_gemm_launches"] and 1)
        launches_per_step = prof["flat_gemm_launches"] / args.steps
        achieved = flops / launches_per_step / t / 1e12
        kk = dim + (3 if metric == "l2" else 0)  # K of the contraction (l2 carries -|x|^2/2 as three bf16 terms)
        nkb = (((kk + 15) // 16 * 16) + 63) // 64
        variant = "flat_gemm_ts_kernel" if nkb <= 4 else ("flat_gemm_tsk_kernel" if nkb <= 12 else "flat_gemm_kernel")
        ridge = B / 2.0  # flop per byte of bf16 X read once
        peak = peaks["bf16_tflops_sustained"]
        bound = "tensor" if ridge >= peak * 1e12 / (peaks["hbm_gbs"] * 1e9) else "hbm"
        if bound == "tensor":
            roof = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                    "traffic": None, "kernel": variant + " (tcgen05 bf16, all segments of a batch)", "peak_source": peaks["source"] + " bf16 sustained"}
        else:
            bytes_ = (n_local * dim * 2.0) / launches_per_step
            a = bytes_ / t / 1e9
            roof = {"bound": "hbm", "achieved": a, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": a / peaks["hbm_gbs"],
                    "traffic": None, "kernel": variant + " (tcgen05 bf16, HBM-bound batch: bf16 shadow read once)", "peak_source": peaks["source"]}
    elif prof["adc_scan_launches"] > 0:
        t = prof["adc_scan_ms"] / prof["adc_scan_launches"] / 1e3
        a = adc_bytes * B / t / 1e9
        roof = {"bound": "hbm", "achieved": a, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": a / peaks["hbm_gbs"],
                "traffic": None, "kernel": "adc_scan_kernel (codes + row terms + ids of the probed lists + LUT)",
                "peak_source": peaks["source"], "avg_launch_ms": t * 1e3, "bytes_per_query_scan": adc_bytes,
                "bytes_per_query_whole_search": ivf_query_bytes,
                "whole_search_gbs": ivf_query_bytes * B * args.steps / (ms / 1e3) / 1e9}
    elif prof["flat_scan_launches"] > 0:
        t = prof["flat_scan_ms"] / prof["flat_scan_launches"] / 1e3
        sel = wl.get("selectivity") or 1.0
        # X read once per pass of <= 4 queries; masked rows are skipped before any load (s*N*D*4 + N/8)
        bytes_ = n_local * dim * 4.0 * sel + (n_local / 8.0 if wl.get("selectivity") else 0.0)
        a = bytes_ / t / 1e9
        roof = {"bound": "hbm", "achieved": a, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": a / peaks["hbm_gbs"],
                "traffic": None, "kernel": "flat_scan_kernel (fp32, <=4 queries per pass)", "peak_source": peaks["source"],
                "launches_per_step": prof["flat_scan_launches"] / args.steps, "avg_launch_ms": t * 1e3,
                "selectivity": sel}

    if roof is not None:
        if roof["bound"] == "hbm" and roof.get("achieved") is not None:
            # SURVEY.md 8d asks for both denominators: the measured copy bandwidth (read + write streams; a read-only
            # scan can exceed it) and the nominal 8 TB/s
            roof["peak_nominal"] = 8000.0
            roof["frac_nominal"] = roof["achieved"] / 8000.0
        # the committed capture is of the default shape of the workload; an overridden shape has no measured traffic
        stock = not args.rows and (args.selectivity is None or args.selectivity == WORKLOADS[args.workload].get("selectivity"))
        roof["traffic"], note = load_traffic(args.workload, B) if stock else (None, None)
        if note:
            roof["traffic_note"] = note
    if rank == 0:
        base = cpu_baseline(wl) if world == 1 and not args.no_cpu_baseline else None
        line = {"metric": "kNN queries/sec", "value": qps, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32" if st1["flat_tc_batches"] == st0["flat_tc_batches"]
                else "bf16 candidates + f32 re-rank", "data": "synthetic N(0,1), seeded", "config": workload_config(wl, args),
                "clocks": clocks,
                "e2e": {"value": qps_e2e, "unit": "queries/s", "h2d_bytes_per_step": int(Qh.numel() * 4 + (allow_host.nbytes if allow_host is not None and world == 1 else 0)),
                        "d2h_bytes_per_step": int(Lh.numel() * 8 + Dh.numel() * 4 + Ch.numel() * 4), "ms_per_step": ms_e2e / args.steps},
                "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": base,
                "paths": {"flat_simt_batches": st1["flat_simt_batches"] - st0["flat_simt_batches"],
                          "flat_tc_batches": st1["flat_tc_batches"] - st0["flat_tc_batches"],
                          "tc_fallback_queries": st1["tc_fallback_queries"] - st0["tc_fallback_queries"],
                          "ivfpq_batches": st1["ivfpq_batches"] - st0["ivfpq_batches"]},
                "profile_ms": {kk: vv for kk, vv in prof.items()}}
        if saved_stdout is not None:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
    h.close()
    if world > 1:
        dist.destroy_process_group()


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
    ap.add_argument("--no-cpu-baseline", action="store_true")
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

"""Row sharding across the GPUs of one box and the merge of the per-shard top-k lists.

One process per GPU.  Rows are split into contiguous label ranges (labels are consecutive,
rust_lib/src/lance_manager.rs:232-233, so a shard is a label range and `label = base + local row`); queries are
replicated; every rank searches its shard; the only exchange is one all_gather of the [nq, k] (label, distance)
lists and the [nq] counts (B*k*12 bytes per rank), after which every rank selects the global top-k.  There is no
data-path collective besides that gather.  The functions take the process group and the final select as
arguments so that the same code runs over NCCL with the device-side select (bench.py) and over gloo on CPU tensors
(tests/test_shard_merge_cpu.py).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(rows: int, rank: int, world: int):
    """[lo, hi) of rank's contiguous row range; sizes differ by at most one row."""
    return rows * rank // world, rows * (rank + 1) // world


class ShardGather:
    """Preallocated gather buffers for one (nq, k) shape.

    Every rank's result is ONE packed record [labels int64 nq*k | dists f32 nq*k | counts i32 nq | overflow i32]
    (padded to 16 B); `L`, `D`, `C`, `O` are views into it, so the search writes its output straight into the record
    and a single all_gather moves it.  `GL`, `GD`, `GC`, `GO` are strided views [world, ...] over the gathered records."""

    def __init__(self, world: int, nq: int, k: int, device, group=None):
        self.world, self.nq, self.k, self.group = world, nq, k, group
        lb, db, cb = nq * k * 8, nq * k * 4, nq * 4
        self.rec_bytes = (lb + db + cb + 4 + 15) // 16 * 16
        self.rec = torch.zeros(self.rec_bytes, dtype=torch.uint8, device=device)
        self.L = self.rec[:lb].view(torch.int64).view(nq, k)
        self.D = self.rec[lb:lb + db].view(torch.float32).view(nq, k)
        self.C = self.rec[lb + db:lb + db + cb].view(torch.int32)
        self.O = self.rec[lb + db + cb:lb + db + cb + 4].view(torch.int32)
        self.all = torch.zeros((world, self.rec_bytes), dtype=torch.uint8, device=device)
        self.GL = self.all[:, :lb].view(torch.int64).view(world, nq, k)
        self.GD = self.all[:, lb:lb + db].view(torch.float32).view(world, nq, k)
        self.GC = self.all[:, lb + db:lb + db + cb].view(torch.int32)
        self.GO = self.all[:, lb + db + cb:lb + db + cb + 4].view(torch.int32)
        self.ML = torch.empty((nq, k), dtype=torch.int64, device=device)
        self.MD = torch.empty((nq, k), dtype=torch.float32, device=device)
        self.MC = torch.empty(nq, dtype=torch.int32, device=device)

    def gather(self, labels: torch.Tensor = None, dists: torch.Tensor = None, counts: torch.Tensor = None):
        """This rank's lists (or, with no arguments, whatever was written into self.L/D/C) -> (GL, GD, GC), strided
        views with the rank as leading dimension."""
        if labels is not None and labels.data_ptr() != self.L.data_ptr():
            self.L.copy_(labels)
            self.D.copy_(dists)
            self.C.copy_(counts)
        if self.world == 1:
            self.all[0].copy_(self.rec)
        else:
            dist.all_gather_into_tensor(self.all.view(-1), self.rec, group=self.group)
        return self.GL, self.GD, self.GC

    def gather_and_merge(self, labels, dists, counts, merge):
        """merge(GL, GD, GC, world, nq, k, ML, MD, MC) performs the final select into the M* buffers."""
        GL, GD, GC = self.gather(labels, dists, counts)
        merge(GL, GD, GC, self.world, self.nq, self.k, self.ML, self.MD, self.MC)
        return self.ML, self.MD, self.MC


class ShardedSearch:
    """One rank's side of a sharded search step: local top-k on this rank's shard -> one all_gather of
    the packed records -> device-side final select, all enqueued on one stream without a host synchronisation
    (lance_b200_search_device_async).  `settle()` then waits for the stream and, if any shard reported an
    overflowed candidate list (degenerate data), has every rank finish its search and repeats gather + select, so
    the merged lists are exact on every rank."""

    def __init__(self, pkg, handle, world: int, nq: int, k: int, device, device_index: int, group=None, merge=None):
        self.pkg, self.h, self.world, self.nq, self.k, self.dev_index = pkg, handle, world, nq, k, device_index
        self.g = ShardGather(world, nq, k, device, group)
        self.merges = 0
        self.stream = 0
        if merge is not None:  # host-side select (CPU tests over gloo)
            self._merge = merge

    def _merge(self, GL, GD, GC, G, nq, k, ML, MD, MC):
        self.pkg.merge_topk_strided_device(self.dev_index, GL.data_ptr(), GD.data_ptr(), GC.data_ptr(), GL.stride(0),
                                           GD.stride(0), GC.stride(0), G, nq, k, ML.data_ptr(), MD.data_ptr(),
                                           MC.data_ptr(), stream=self.stream)
        self.merges += 1

    def step(self, d_queries: int, stream: int, **search_kw):
        g = self.g
        self.stream = stream
        self.h.search_device_async(d_queries, self.nq, self.k, g.L.data_ptr(), g.D.data_ptr(), g.C.data_ptr(),
                                   stream=stream, d_overflow=g.O.data_ptr(), **search_kw)
        return g.gather_and_merge(None, None, None, self._merge)

    def settle(self) -> int:
        """Host-side end of a step; returns how many queries any shard had to re-run (0 on non-degenerate data)."""
        g = self.g
        pending = int(g.GO.sum().item())  # waits for the stream
        if pending == 0:
            return 0
        self.h.search_finish(self.stream)
        g.O.zero_()
        g.gather_and_merge(None, None, None, self._merge)
        if g.rec.is_cuda:
            torch.cuda.current_stream().synchronize()
        return pending


# ---------------------------------------------------------------------------------------------------------------------
# IVF_PQ: shard BY PARTITION.  One global index (centroids and codebooks replicated, 13.4 MB at C3); the
# inverted lists are dealt to the ranks whole, balanced by length, so every probed list is scanned on exactly one GPU.
# Every rank probes the SAME nprobes partitions (same centroids, same arithmetic), scans the ones it owns (the others
# are empty lists in its copy of the index) and refines with its own rows; the union of the ranks' candidates is the
# one-GPU candidate set, so the merged top-k equals the one-GPU result.
def assign_partitions(part_sizes, world: int):
    """Longest-processing-time deal of the partitions to `world` ranks: -> owner[P] (int array)."""
    import numpy as np

    sizes = np.asarray(part_sizes, np.int64)
    owner = np.zeros(sizes.shape[0], np.int32)
    load = np.zeros(world, np.int64)
    for p in np.argsort(-sizes, kind="stable"):
        r = int(np.argmin(load))
        owner[p] = r
        load[r] += sizes[p]
    return owner


def ivfpq_partition_shard(pkg, index: dict, X_dev, rank: int, world: int, metric: str, stream: int = 0):
    """Builds rank's shard handle from a global index in the exchange layout (`LanceHandle.get_ivfpq()` of a handle whose
    label base is 0) and the device-resident table X_dev [rows, dim] (a torch tensor on this rank's device).
    Returns (handle, owner): the handle holds only the rows of the partitions this rank owns, under their global
    labels, and an index whose other partitions are empty."""
    import numpy as np
    import torch

    off, ids, codes = index["part_offsets"], index["row_ids"], index["codes"]
    P, dim = index["P"], index["dim"]
    owner = assign_partitions(np.diff(off), world)
    mine = np.flatnonzero(owner == rank)
    seg = [np.arange(off[p], off[p + 1]) for p in mine]
    pos = np.concatenate(seg) if seg else np.zeros(0, np.int64)  # positions in the exchange arrays, partition-major
    labels = ids[pos]
    order = np.argsort(labels, kind="stable")  # rows of the shard: ascending label
    row_of_pos = np.empty_like(order)
    row_of_pos[order] = np.arange(order.shape[0])
    h = pkg.LanceHandle(dim, metric)
    sorted_labels = labels[order]
    if sorted_labels.shape[0]:
        sel = torch.from_numpy(sorted_labels).to(X_dev.device)
        step = 1 << 20
        for lo in range(0, sel.shape[0], step):  # gather on the device, append with the global labels
            rows = X_dev.index_select(0, sel[lo:lo + step]).contiguous()
            h.add_batch_device_labels(rows.data_ptr(), rows.shape[0], dim, sorted_labels[lo:lo + step], stream)
    sizes = np.where(owner == rank, np.diff(off), 0)
    new_off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    # explicit labels: the exchange layout is in ROW numbers of the shard
    h.set_ivfpq(index["centroids"], index["codebook"], new_off, row_of_pos.astype(np.int64), codes[pos])
    return h, owner

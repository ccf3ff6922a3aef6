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

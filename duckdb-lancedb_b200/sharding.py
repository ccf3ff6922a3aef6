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
    """Preallocated gather buffers for one (nq, k) shape."""

    def __init__(self, world: int, nq: int, k: int, device, group=None):
        self.world, self.nq, self.k, self.group = world, nq, k, group
        self.GL = torch.empty((world, nq, k), dtype=torch.int64, device=device)
        self.GD = torch.empty((world, nq, k), dtype=torch.float32, device=device)
        self.GC = torch.empty((world, nq), dtype=torch.int32, device=device)
        self.ML = torch.empty((nq, k), dtype=torch.int64, device=device)
        self.MD = torch.empty((nq, k), dtype=torch.float32, device=device)
        self.MC = torch.empty(nq, dtype=torch.int32, device=device)

    def gather(self, labels: torch.Tensor, dists: torch.Tensor, counts: torch.Tensor):
        """labels/dists [nq, k], counts [nq] of this rank -> (GL, GD, GC) with the rank as leading dimension."""
        if self.world == 1:
            self.GL[0].copy_(labels)
            self.GD[0].copy_(dists)
            self.GC[0].copy_(counts)
        else:
            # output viewed as the concatenation along dim 0 (the form both NCCL and gloo accept)
            w, nq, k = self.world, self.nq, self.k
            dist.all_gather_into_tensor(self.GL.view(w * nq, k), labels.contiguous(), group=self.group)
            dist.all_gather_into_tensor(self.GD.view(w * nq, k), dists.contiguous(), group=self.group)
            dist.all_gather_into_tensor(self.GC.view(w * nq), counts.contiguous(), group=self.group)
        return self.GL, self.GD, self.GC

    def gather_and_merge(self, labels, dists, counts, merge):
        """merge(GL, GD, GC, world, nq, k, ML, MD, MC) performs the final select into the M* buffers."""
        GL, GD, GC = self.gather(labels, dists, counts)
        merge(GL, GD, GC, self.world, self.nq, self.k, self.ML, self.MD, self.MC)
        return self.ML, self.MD, self.MC

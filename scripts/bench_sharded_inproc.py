#!/usr/bin/env python
"""bench_sharded_inproc.py [rows_total] [batch] — the in-process sharded handle (lance_b200_create_sharded) against a
plain one-device handle on the same table: host-buffer lance_detached_search_batch calls, wall-clock per call."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
d, k = 128, 10
G = pkg.device_count()
rng = np.random.default_rng(1)
Q = rng.standard_normal((B, d)).astype(np.float32)
out = {"rows": rows, "dim": d, "batch": B, "k": k, "devices": G, "metric": "dot"}
import torch  # device memory for the generator only: the rows go in through the host-buffer C call

handles = (("one_device", pkg.LanceHandle(d, "dot")), (f"sharded_{G}", pkg.LanceHandle.create_sharded(list(range(G)), d, "dot")))
gen = torch.Generator(device="cuda:0")
gen.manual_seed(2)
for lo in range(0, rows, 1 << 20):
    chunk = torch.randn((min(1 << 20, rows - lo), d), generator=gen, device="cuda:0", dtype=torch.float32).cpu().numpy()
    for _, handle in handles:
        handle.add_batch(chunk)
for name, handle in handles:
    res = None
    for _ in range(3):
        res = handle.search_batch(Q, k)
    t0 = time.perf_counter()
    steps = 10
    for _ in range(steps):
        res = handle.search_batch(Q, k)
    ms = (time.perf_counter() - t0) / steps * 1e3
    out[name] = {"ms_per_batch": ms, "queries_per_s": B / ms * 1e3, "labels_checksum": int(res[0].sum())}
    handle.close()
out["speedup"] = out["one_device"]["ms_per_batch"] / out[f"sharded_{G}"]["ms_per_batch"]
out["same_results"] = out["one_device"]["labels_checksum"] == out[f"sharded_{G}"]["labels_checksum"]
print(json.dumps(out))

import sys, time, os
sys.path.insert(0, "/root/repo")
from __graft_entry__ import load_package
import torch, numpy as np
pkg = load_package()
n, d = 4_000_000, 768
h = pkg.LanceHandle(d, "l2")
g = torch.Generator(device="cuda:0"); g.manual_seed(3)
for lo in range(0, n, 1 << 19):
    x = torch.randn((min(1 << 19, n - lo), d), generator=g, device="cuda:0")
    h.add_batch_device(x.data_ptr(), x.shape[0], d)
torch.cuda.synchronize()
for tc in ("0", "1"):
    os.environ["LANCE_B200_TRAIN_TC"] = tc
    t0 = time.perf_counter(); h.create_index(2048, 96); dt = time.perf_counter() - t0
    Q = torch.randn((64, d), generator=g, device="cuda:0").cpu().numpy()
    L, D, C = h.search_batch(Q, 10, nprobes=32, refine_factor=2)
    h.drop_index()
    print("TRAIN_TC=%s: create_index %.2f s, first labels %s" % (tc, dt, L[0, :4].tolist()))
h.close()

"""Development tool: what the PCIe link of this box gives for one large pinned copy in either
direction (torch, one cudaMemcpyAsync each), next to b200lb_f_upload / _download of the same
bytes (staged in 64 MB pieces with the layout conversion on the device)."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ludwig_b200 as lb                       # noqa: E402
from ludwig_b200 import models_host, synthetic  # noqa: E402

n = 258 * 258 * 258 * 19
host = torch.empty(n, dtype=torch.float64).pin_memory()
dev = torch.empty(n, dtype=torch.float64, device="cuda")
out = {}
for name, fn in (("h2d", lambda: dev.copy_(host, non_blocking=True)),
                 ("d2h", lambda: host.copy_(dev, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
    out["raw_%s_GBps" % name] = n * 8 / best / 1e9
del dev
local = (256, 256, 256)
lat = lb.Lattice(19, local)
f0 = synthetic.make_state(models_host.cv(19), models_host.wv(19), local)
hbuf = lb.pinned_empty(f0.shape)
hbuf[...] = f0
for name, fn in (("h2d", lambda: lat.lb_memcpy_h2d(hbuf)), ("d2h", lambda: lat.lb_memcpy_d2h(hbuf))):
    fn(); lat.sync()
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter(); fn(); lat.sync(); best = min(best, time.perf_counter() - t0)
    out["lb_memcpy_%s_GBps" % name] = hbuf.nbytes / best / 1e9
lat.close()
print(json.dumps(out))

#!/usr/bin/env python
"""Small driver for profiling stage 3 on the C2 cloud: builds the map once, then runs a few pm_knn
passes on device-resident buffers. Usage: python tools/prof_knn.py [n_photons] [n_queries] [k] [steps]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402
import bench  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
k = int(sys.argv[3]) if len(sys.argv) > 3 else 64
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
pkg = graft.load_package()
ctx = pkg.Context(0)
cloud = bench.make_cloud(n)
d_tree = torch.from_numpy(cloud.view(np.uint8)).cuda()
torch.cuda.synchronize()
t0 = time.perf_counter()
pmap = pkg.PhotonMap.build(ctx, d_tree)
print(f"build {1e3 * (time.perf_counter() - t0):.2f} ms", flush=True)
q = bench.make_queries(nq, 0)
if n != 10_000_000:
    q *= np.float32(n * bench.SCALE / bench.QUERY_EXTENT)
d_q = torch.from_numpy(q).cuda()
d_ids = torch.empty((nq, k), dtype=torch.int32, device="cuda")
d_d2 = torch.empty((nq, k), dtype=torch.float32, device="cuda")
d_r2 = torch.empty(nq, dtype=torch.float32, device="cuda")
for i in range(steps):
    t0 = time.perf_counter()
    pmap.knn(d_q, k, 100.0, out_ids=d_ids, out_dist2=d_d2, out_r2=d_r2)
    print(f"step {i}: call {1e3 * (time.perf_counter() - t0):.3f} ms, query kernels {ctx.last_kernel_ms:.3f} ms, "
          f"kd-tree fallbacks {ctx.last_fallback_count}", flush=True)

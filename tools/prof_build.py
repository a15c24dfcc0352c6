#!/usr/bin/env python
"""Builds the C2-style map a few times (for launch lists). Usage: python tools/prof_build.py [n_photons] [reps]"""
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
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
pkg = graft.load_package()
ctx = pkg.Context(0)
cloud = bench.make_cloud(n)
for r in range(reps):
    d_tree = torch.from_numpy(cloud.view(np.uint8)).cuda()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    pmap = pkg.PhotonMap.build(ctx, d_tree)
    print(f"build {r}: {1e3 * (time.perf_counter() - t0):.2f} ms", flush=True)
    del pmap

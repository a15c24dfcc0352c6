#!/usr/bin/env python
"""Times warm pm_map_build calls on the C2 cloud: min / median of N builds. Usage: python tools/time_build.py [n_photons] [reps] [carry_from]"""
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
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 12
pkg = graft.load_package()
if len(sys.argv) > 3:  # force the carried-keys variant of the level pass from this many photons on
    pkg.lib().pm_test_build_carry_from(int(sys.argv[3]))
ctx = pkg.Context(0)
cloud = torch.from_numpy(bench.make_cloud(n).view(np.uint8))
d_tree = cloud.cuda()
ms = []
for r in range(reps + 2):
    d_tree.copy_(cloud)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    pmap = pkg.PhotonMap.build(ctx, d_tree)
    ms.append(1e3 * (time.perf_counter() - t0))
    pmap.close()
ms = sorted(ms[2:])
print(f"build n={n}: min {ms[0]:.3f} ms, median {ms[len(ms) // 2]:.3f} ms")

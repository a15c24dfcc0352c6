#!/usr/bin/env python
"""Times pm_map_build_shard for emulated ranks on one GPU against the whole build (n random photons).
Usage: python tools/prof_shard_build.py [n]"""
import ctypes
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
pkg = graft.load_package()
ctx = pkg.Context(0)
L = pkg.lib()
g = torch.Generator(device="cuda").manual_seed(3)
ph = torch.zeros((n, 11), dtype=torch.float32, device="cuda")
ph[:, 0:3] = torch.rand((n, 3), generator=g, device="cuda") * 40.0
ph[:, 6:10] = 1.0
src = ph.clone()


def timed(f):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = f()
    torch.cuda.synchronize()
    return r, 1e3 * (time.perf_counter() - t0)


for it in range(2):
    ph.copy_(src)
    m, ms = timed(lambda: pkg.PhotonMap.build(ctx, ph))
    m.close()
print(f"pm_map_build {n}: {ms:.1f} ms", flush=True)
elem = torch.empty(n, dtype=torch.int32, device="cuda")
dim = torch.zeros((n + 3) // 4 * 4, dtype=torch.uint8, device="cuda")
bounds = np.zeros(6, np.float32)
ph.copy_(src)
for world in (1, 2, 8):
    for rank in sorted({0, world - 1}):
        for it in range(2):
            _, ms = timed(lambda: pkg._check(L.pm_map_build_shard(ctx.handle, ph.data_ptr(), n, rank, world, elem.data_ptr(),
                                                                   dim.data_ptr(), bounds.ctypes.data)))
        print(f"pm_map_build_shard rank {rank} of {world}: {ms:.1f} ms", flush=True)
_ = pkg._check(L.pm_map_build_shard(ctx.handle, ph.data_ptr(), n, 0, 1, elem.data_ptr(), dim.data_ptr(), bounds.ctypes.data))
h = ctypes.c_void_p()
_, ms = timed(lambda: pkg._check(L.pm_map_build_finish(ctx.handle, ph.data_ptr(), n, elem.data_ptr(), dim.data_ptr(), bounds.ctypes.data,
                                                       ctypes.byref(h))))
print(f"pm_map_build_finish: {ms:.1f} ms", flush=True)

#!/usr/bin/env python
"""Stage-1 profiling driver: Cornell box (58 triangles) or the procedural 1 M-triangle scene, device-resident output.
Usage: python tools/prof_emit.py [cornell|big] [casted] [reps]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "cornell"
casted = int(sys.argv[2]) if len(sys.argv) > 2 else 4_000_000
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
pkg = graft.load_package()
ctx = pkg.Context(0)
if which == "cornell":
    z = np.load(os.path.join(ROOT, "tests", "golden", "scene_cornell_box.npz"))
    scene = pkg.Scene(ctx, z["verts"], z["tris"], z["tri_mesh"], z["mats"].view(pkg.MATERIAL).reshape(-1))
    lights = z["lights"].view(pkg.LIGHT).reshape(-1)
else:
    # the C4 room (photon-mapping_b200/scenes.py): displaced sphere of 1 M triangles inside a closed, inward-facing box
    sc = graft.load_package_module("scenes").c4_room()
    scene = pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])
    lights = np.ascontiguousarray(sc["lights"], pkg.LIGHT)
depth = 8
cap = casted * depth
d_out = torch.empty(cap * 40, dtype=torch.uint8, device="cuda")
for it in range(reps):
    t0 = time.perf_counter()
    _, n, paths, segs = scene.emit_photons(lights, casted, depth, False, seed=1, rng_mode=pkg.RNG_COUNTER, out=d_out, capacity=cap)
    wall = time.perf_counter() - t0
    print(f"rep {it}: call {1e3 * wall:.3f} ms, walk kernel {ctx.last_kernel_ms:.3f} ms, paths {paths}, segments {segs}, stored {n}",
          flush=True)

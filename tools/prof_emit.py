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
    # the C4-like scene of tools/scale_check.py: displaced sphere of 1 M triangles inside a closed box
    nu, nv = 500, 1000
    u = np.linspace(0, np.pi, nu + 1, dtype=np.float64)[:, None]
    v = np.linspace(0, 2 * np.pi, nv + 1, dtype=np.float64)[None, :]
    r = 12.0 * (1 + 0.08 * np.sin(7 * u) * np.sin(5 * v) + 0.04 * np.sin(23 * u + 1.3) * np.sin(17 * v) +
                0.02 * np.sin(61 * u) * np.sin(53 * v + 0.7) + 0.01 * np.sin(131 * u + 2.1) * np.sin(113 * v))
    sv = (np.stack([r * np.sin(u) * np.cos(v), r * np.cos(u) + 0 * v, r * np.sin(u) * np.sin(v)], -1) + [0.0, 14.0, 0.0])
    sv = sv.reshape(-1, 3).astype(np.float32)
    idx = np.arange((nu + 1) * (nv + 1)).reshape(nu + 1, nv + 1)
    a, b, c, d = idx[:-1, :-1], idx[1:, :-1], idx[1:, 1:], idx[:-1, 1:]
    st = np.concatenate([np.stack([a, b, c], -1).reshape(-1, 3), np.stack([a, c, d], -1).reshape(-1, 3)]).astype(np.int32)
    lo, hi = np.array([-20, 0, -20], np.float32), np.array([20, 40, 20], np.float32)
    bv = np.array([[x, y, z] for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])], np.float32)
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    bt = np.array([t for q in quads for t in ((q[0], q[1], q[2]), (q[0], q[2], q[3]))], np.int32)
    verts, tris = np.concatenate([sv, bv]), np.concatenate([st, bt + len(sv)])
    tri_mesh = np.concatenate([np.zeros(len(st), np.int32), 1 + (np.arange(len(bt)) // 2).astype(np.int32)])
    mats = np.zeros(7, pkg.MATERIAL)
    mats[0] = ((0.8, 0.7, 0.6), 0.7, 0.2, 0.0, 1.0)
    for i in range(1, 7):
        mats[i] = ((0.75, 0.75, 0.75) if i > 2 else ((0.8, 0.2, 0.2) if i == 1 else (0.2, 0.8, 0.2)), 0.9, 0.0, 0.0, 1.0)
    lights = np.zeros(2, pkg.LIGHT)
    lights[0] = ((-12.0, 36.0, -8.0), (1, 1, 1), 10.0)
    lights[1] = ((13.0, 35.0, 11.0), (1.0, 0.9, 0.8), 10.0)
    scene = pkg.Scene(ctx, verts, tris, tri_mesh, mats)
depth = 8
cap = casted * depth
d_out = torch.empty(cap * 40, dtype=torch.uint8, device="cuda")
for it in range(reps):
    t0 = time.perf_counter()
    _, n, paths, segs = scene.emit_photons(lights, casted, depth, False, seed=1, rng_mode=pkg.RNG_COUNTER, out=d_out, capacity=cap)
    wall = time.perf_counter() - t0
    print(f"rep {it}: call {1e3 * wall:.3f} ms, walk kernel {ctx.last_kernel_ms:.3f} ms, paths {paths}, segments {segs}, stored {n}",
          flush=True)

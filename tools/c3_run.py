#!/usr/bin/env python
"""C3-like end-to-end pass on one B200, everything device resident: Cornell box, N cast photons (+ N/10 caustic),
max_depth 8, counter RNG -> maps -> 1920x1080 render (spp 1, depth 1). Writes gpurun_out/c3_run.json.
Usage: python tools/c3_run.py [casted] [W] [H]"""
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

casted = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
H = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
pkg = graft.load_package()
ctx = pkg.Context(0)
z = np.load(os.path.join(ROOT, "tests", "golden", "scene_cornell_box.npz"))
scene = pkg.Scene(ctx, z["verts"], z["tris"], z["tri_mesh"], z["mats"].view(pkg.MATERIAL).reshape(-1))
lights = z["lights"].view(pkg.LIGHT).reshape(-1)
res = {"casted": casted, "fb": [W, H]}


def timed(f):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = f()
    torch.cuda.synchronize()
    return r, 1e3 * (time.perf_counter() - t0)


depth = 8
cap_g, cap_c = casted * 3, max(casted // 10, 1) * 2
d_g = torch.empty(cap_g * 40, dtype=torch.uint8, device="cuda")
d_c = torch.empty(cap_c * 40, dtype=torch.uint8, device="cuda")
(_, ng, paths, segs), ms_first = timed(lambda: scene.emit_photons(lights, casted, depth, False, seed=1, rng_mode=pkg.RNG_COUNTER, out=d_g, capacity=cap_g))
(_, ng, paths, segs), ms = timed(lambda: scene.emit_photons(lights, casted, depth, False, seed=1, rng_mode=pkg.RNG_COUNTER, out=d_g, capacity=cap_g))
res["emit_global"] = {"ms": ms, "ms_first": ms_first, "paths": paths, "segments": segs, "stored": ng, "walk_kernel_ms": ctx.last_kernel_ms}
(_, nc, paths_c, segs_c), ms = timed(lambda: scene.emit_photons(lights, casted // 10, depth, True, seed=1, rng_mode=pkg.RNG_COUNTER, out=d_c, capacity=cap_c))
res["emit_caustic"] = {"ms": ms, "paths": paths_c, "segments": segs_c, "stored": nc}
g44 = torch.empty((ng + nc) * 44, dtype=torch.uint8, device="cuda")
c44 = torch.empty(max(nc, 1) * 44, dtype=torch.uint8, device="cuda")
_, ms = timed(lambda: pkg._check(pkg.lib().pm_photons_to_maps(ctx.handle, d_g.data_ptr(), ng, d_c.data_ptr() if nc else None, nc,
                                                               g44.data_ptr(), c44.data_ptr() if nc else None)))
res["to_maps_ms"] = ms
del d_g, d_c
g_copy, c_copy = g44.clone(), c44.clone()
gmap, ms0 = timed(lambda: pkg.PhotonMap.build(ctx, g_copy))  # first build grows the context's scratch buffers
del gmap, g_copy
gmap, ms = timed(lambda: pkg.PhotonMap.build(ctx, g44))
res["build_global"] = {"photons": ng + nc, "ms": ms, "ms_first": ms0}
cmap, ms0 = timed(lambda: pkg.PhotonMap.build(ctx, c_copy[: nc * 44]) if nc else None)
del cmap, c_copy
cmap, ms = timed(lambda: pkg.PhotonMap.build(ctx, c44[: nc * 44]) if nc else None)
res["build_caustic"] = {"photons": nc, "ms": ms, "ms_first": ms0}
cam = pkg.make_camera((80.0, 30.0, 0.0), (10.0, 20.0, 0.0), (0.0, 1.0, 0.0), 0.87, W, H)
for it in range(2):
    fb, ms = timed(lambda: scene.render(lights, gmap, cmap, cam, W, H, 1, 1))
res["render"] = {"ms": ms, "gather_kernels_ms_last_wave": ctx.last_kernel_ms, "lit_fraction": float((fb[1:] != 0).mean()),
                 "kd_fallbacks": ctx.last_fallback_count}
# gather microbench on the real (surface) cloud: 2 M hit points = positions of a strided sample of the photons
nq = 2_000_000
pos = g44.view(-1, 44)[:: max((ng + nc) // nq, 1)][:nq, :12].contiguous().view(torch.float32).view(-1, 3).contiguous()
brdf = torch.full((len(pos),), 0.25, device="cuda")
for it in range(2):
    _, ms = timed(lambda: gmap.gather(pos, brdf, 50, 100.0))
res["gather_on_surface_cloud"] = {"queries": int(len(pos)), "ms": ms, "kernels_ms": ctx.last_kernel_ms,
                                  "kd_fallbacks": ctx.last_fallback_count, "handovers": ctx.last_handover_count}
res["total_ms"] = res["emit_global"]["ms"] + res["emit_caustic"]["ms"] + res["to_maps_ms"] + res["build_global"]["ms"] + \
    res["build_caustic"]["ms"] + res["render"]["ms"]
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "c3_run.json"), "w"), indent=1)
print(json.dumps(res))

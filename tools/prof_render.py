#!/usr/bin/env python
"""C1-like render for profiling the gather kernels on a surface-like cloud: Cornell box, casted photons (default
2 M + 200 k), W x H (default 1024 x 1024), spp 1, depth 1.  usage: python tools/prof_render.py [casted] [W] [H] [reps]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft
casted = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
H = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
pkg = graft.load_package()
ctx = pkg.Context(0)
z = np.load(os.path.join(ROOT, "tests", "golden", "scene_cornell_box.npz"))
mats = z["mats"].view(pkg.MATERIAL).reshape(-1); lights = z["lights"].view(pkg.LIGHT).reshape(-1)
scene = pkg.Scene(ctx, z["verts"], z["tris"], z["tri_mesh"], mats)
g40, ng, _, _ = scene.emit_photons(lights, casted, 8, False, seed=1, rng_mode=pkg.RNG_COUNTER)
c40, nc, _, _ = scene.emit_photons(lights, casted // 10, 8, True, seed=1, rng_mode=pkg.RNG_COUNTER)
g44, c44 = pkg.photons_to_maps(ctx, g40, c40)
gmap = pkg.PhotonMap.build(ctx, g44); cmap = pkg.PhotonMap.build(ctx, c44)
cam = pkg.make_camera((80.0, 30.0, 0.0), (10.0, 20.0, 0.0), (0.0, 1.0, 0.0), 0.87, W, H)
for it in range(reps):
    t0 = time.perf_counter()
    fb = scene.render(lights, gmap, cmap, cam, W, H, 1, 1)
    print(f"render {it}: {1e3 * (time.perf_counter() - t0):.2f} ms, global-map gather kernels {ctx.last_kernel_ms:.2f} ms, "
          f"photons {len(g44)} + {len(c44)}, kd fallbacks {ctx.last_fallback_count}, lit {float((fb[1:] != 0).mean()):.3f}", flush=True)
print("hand-overs of the last gather call:", ctx.last_handover_count, "kd fallbacks:", ctx.last_fallback_count)

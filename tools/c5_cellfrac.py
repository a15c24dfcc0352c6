#!/usr/bin/env python
"""C5 gather time on ONE GPU for the photon share one of `ranks` ranks would hold, for several cell fractions.
usage: python tools/c5_cellfrac.py [ranks]"""
import ctypes, os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft
ranks = int(sys.argv[1]) if len(sys.argv) > 1 else 8
pkg = graft.load_package(); ppm = graft.load_package_module("ppm")
ctx = pkg.Context(0); L = pkg.lib()
z = np.load(os.path.join(ROOT, "tests", "golden", "scene_cornell_box.npz"))
mats = z["mats"].view(pkg.MATERIAL).reshape(-1); lights = np.ascontiguousarray(z["lights"].view(pkg.LIGHT).reshape(-1), pkg.LIGHT)
scene = pkg.Scene(ctx, z["verts"], z["tris"], z["tri_mesh"], mats)
W, H = 3840, 2160
cam = pkg.make_camera((80.0, 30.0, 0.0), (10.0, 20.0, 0.0), (0.0, 1.0, 0.0), 0.87, W, H)
rays = ppm.camera_rays(pkg, cam, W, H)
pos_h, brdf_h, _ = ppm.hit_points(pkg, scene, np.asarray(z["tri_mesh"]), mats, rays)
dev = torch.device("cuda", 0)
pos, brdf = torch.from_numpy(pos_h).to(dev), torch.from_numpy(brdf_h).to(dev)
npx = W * H
per_pass, depth = 15_625_000, 8
cap = (per_pass // ranks + 8) * depth
d40 = torch.empty((cap, 40), dtype=torch.uint8, device=dev); g44 = torch.empty((cap, 44), dtype=torch.uint8, device=dev)
flux = torch.empty((npx, 3), dtype=torch.float32, device=dev); count = torch.empty(npx, dtype=torch.int32, device=dev)
cnt, npaths, nseg = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int64(0)
pkg._check(L.pm_emit_photons(ctx.handle, scene.handle, lights.ctypes.data, len(lights), per_pass, depth, 0, 1, pkg.RNG_COUNTER, 0, ranks,
                             d40.data_ptr(), cap, ctypes.byref(cnt), ctypes.byref(npaths), ctypes.byref(nseg)))
n = cnt.value
pkg._check(L.pm_photons_to_maps(ctx.handle, d40.data_ptr(), n, None, 0, g44.data_ptr(), None))
fracs = [float(x) for x in os.environ.get("C5_FRACS", "0.15,0.2,0.3,0.4,0.5,0.7,1.0").split(",")]
for r in (0.5, 0.35, 0.25):
    r2 = torch.full((npx,), r * r, dtype=torch.float32, device=dev)
    for frac in fracs:
        ms = []
        for it in range(3):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            pmap = pkg.PhotonMap.build_cells(ctx, g44[:n], frac * r)
            torch.cuda.synchronize(); t1 = time.perf_counter()
            pmap.gather_radius(pos, brdf, r2, out_flux=flux, out_count=count)
            torch.cuda.synchronize(); t2 = time.perf_counter()
            ms.append((1e3 * (t1 - t0), 1e3 * (t2 - t1), ctx.last_kernel_ms))
            pmap.close()
        b = min(ms[1:], key=lambda x: x[0] + x[1])
        print(f"ranks {ranks} photons {n} radius {r} cell_frac {frac}: map {b[0]:.2f} ms gather {b[1]:.2f} ms (kernels {b[2]:.2f}) heavy {ctx.last_handover_count} mean count {float(count.float().mean()):.1f}", flush=True)

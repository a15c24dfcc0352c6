#!/usr/bin/env python
"""Scale check at BASELINE.json's larger configurations on one B200 (no oracle: properties only).
  C4-like: procedural 1 M-triangle scene (closed box + displaced sphere), 25 M cast photons (one rank's share of
           200 M / 8), counter RNG, max_depth 8 -> photons/s; hit ids of 100 k rays are cross-checked between the
           closest-hit kernel and an occlusion query.
  C3-like: 50 M-photon map build + 2 M fused gathers (k = 50).
Writes gpurun_out/scale_check.json."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

pkg = graft.load_package()
ctx = pkg.Context(0)
res = {}


def displaced_sphere(nu, nv, radius, centre):
    u = np.linspace(0, np.pi, nu + 1, dtype=np.float64)[:, None]
    v = np.linspace(0, 2 * np.pi, nv + 1, dtype=np.float64)[None, :]
    r = radius * (1 + 0.08 * np.sin(7 * u) * np.sin(5 * v) + 0.04 * np.sin(23 * u + 1.3) * np.sin(17 * v) +
                  0.02 * np.sin(61 * u) * np.sin(53 * v + 0.7) + 0.01 * np.sin(131 * u + 2.1) * np.sin(113 * v))
    p = np.stack([r * np.sin(u) * np.cos(v), r * np.cos(u) + 0 * v, r * np.sin(u) * np.sin(v)], -1) + np.asarray(centre)
    idx = np.arange((nu + 1) * (nv + 1)).reshape(nu + 1, nv + 1)
    a, b, c, d = idx[:-1, :-1], idx[1:, :-1], idx[1:, 1:], idx[:-1, 1:]
    tris = np.concatenate([np.stack([a, b, c], -1).reshape(-1, 3), np.stack([a, c, d], -1).reshape(-1, 3)])
    return p.reshape(-1, 3).astype(np.float32), tris.astype(np.int32)


def box(lo, hi):
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    c = np.array([[x, y, z] for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])], np.float32)
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    tris = np.array([t for q in quads for t in ((q[0], q[1], q[2]), (q[0], q[2], q[3]))], np.int32)
    return c, tris


# ---- C4-like scene ---------------------------------------------------------------------------------
sv, st = displaced_sphere(500, 1000, 12.0, (0.0, 14.0, 0.0))  # 1 000 000 triangles
bv, bt = box((-20, 0, -20), (20, 40, 20))
verts = np.concatenate([sv, bv])
tris = np.concatenate([st, bt + len(sv)])
tri_mesh = np.concatenate([np.zeros(len(st), np.int32), 1 + (np.arange(len(bt)) // 2).astype(np.int32)])
mats = np.zeros(7, pkg.MATERIAL)
mats[0] = ((0.8, 0.7, 0.6), 0.7, 0.2, 0.0, 1.0)
for i in range(1, 7):
    mats[i] = ((0.75, 0.75, 0.75) if i > 2 else ((0.8, 0.2, 0.2) if i == 1 else (0.2, 0.8, 0.2)), 0.9, 0.0, 0.0, 1.0)
lights = np.zeros(2, pkg.LIGHT)
lights[0] = ((-12.0, 36.0, -8.0), (1, 1, 1), 10.0)
lights[1] = ((13.0, 35.0, 11.0), (1.0, 0.9, 0.8), 10.0)
t0 = time.perf_counter()
scene = pkg.Scene(ctx, verts, tris, tri_mesh, mats)
res["c4_scene"] = {"triangles": int(len(tris)), "scene_create_ms": 1e3 * (time.perf_counter() - t0)}

rng = np.random.default_rng(1)
rays = np.zeros(100_000, pkg.RAY)
rays["org"] = (rng.random((len(rays), 3)) * [36, 36, 36] + [-18, 2, -18]).astype(np.float32)
d = rng.standard_normal((len(rays), 3)).astype(np.float32)
rays["dir"] = d / np.linalg.norm(d, axis=1, keepdims=True)
rays["tmin"], rays["tmax"] = 1e-3, 1e30
hits = scene.trace_closest(rays)
res["c4_scene"]["trace_100k_rays_kernel_ms"] = ctx.last_kernel_ms
assert (hits["prim"] >= 0).mean() > 0.99  # closed box
sh = rays.copy()
sh["tmax"] = np.where(hits["prim"] >= 0, hits["t"] * 0.999, 1e30).astype(np.float32)
assert scene.trace_occluded(sh).sum() == 0  # nothing in front of the closest hit
sh["tmax"] = np.where(hits["prim"] >= 0, hits["t"] * 1.001 + 1e-3, 1e30).astype(np.float32)
assert scene.trace_occluded(sh).mean() > 0.99

casted, depth = 25_000_000, 8
cap = casted * 3
d_out = torch.empty(cap * 40, dtype=torch.uint8, device="cuda")
best = None
for it in range(2):
    t0 = time.perf_counter()
    _, n, paths, segs = scene.emit_photons(lights, casted, depth, False, seed=1, rng_mode=pkg.RNG_COUNTER, out=d_out, capacity=cap)
    wall = time.perf_counter() - t0
    best = (wall, ctx.last_kernel_ms, n, paths, segs)
res["c4_emit"] = {"casted": casted, "paths": best[3], "segments": best[4], "stored": best[2], "call_ms": 1e3 * best[0],
                  "walk_kernel_ms": best[1], "paths_per_s": best[3] / best[0], "segments_per_s": best[4] / best[0]}
pos = d_out[: best[2] * 40].view(-1, 40)[:, :12].contiguous().view(torch.float32).view(-1, 3)
assert bool(torch.isfinite(pos).all()) and float(pos[:, 1].min()) > -0.01 and float(pos[:, 1].max()) < 40.01
del d_out, pos
scene.close()

# ---- C3-like map ------------------------------------------------------------------------------------
n = 50_000_000
g = torch.Generator(device="cuda").manual_seed(3)
ph = torch.zeros((n, 11), dtype=torch.float32, device="cuda")
ph[:, 0:3] = torch.rand((n, 3), generator=g, device="cuda") * torch.tensor([40.0, 40.0, 40.0], device="cuda")
ph[:, 6:9] = 1.0
ph[:, 9] = 1.0
torch.cuda.synchronize()
t0 = time.perf_counter()
pmap = pkg.PhotonMap.build(ctx, ph)
build_ms = 1e3 * (time.perf_counter() - t0)
nq = 2_000_000
q = torch.rand((nq, 3), generator=g, device="cuda") * 40.0
brdf = torch.full((nq,), 0.25, device="cuda")
rgb = pmap.gather(q, brdf, 50, 100.0)
rgb = pmap.gather(q, brdf, 50, 100.0)
res["c3_map"] = {"photons": n, "build_ms": build_ms, "build_photons_per_s": n / (build_ms * 1e-3), "gather_queries": nq, "k": 50,
                 "gather_kernels_ms": ctx.last_kernel_ms, "gather_queries_per_s": nq / (ctx.last_kernel_ms * 1e-3),
                 "kd_tree_fallback_queries": ctx.last_fallback_count}
# uniform density rho, power 1, colour 1: estimate ~ brdf * k * mean(w) / (c * r^2) is positive and tightly distributed
assert bool(torch.isfinite(rgb).all()) and float(rgb.min()) > 0
ids, d2, r2 = pmap.knn(q[:100000].contiguous(), 64, 100.0)
assert bool((ids >= 0).all()) and bool((d2.max(dim=1).values == r2).all())
res["c3_map"]["rgb_mean"] = float(rgb.mean())
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "scale_check.json"), "w"), indent=1)
print(json.dumps(res))

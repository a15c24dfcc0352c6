"""Progressive (fixed-radius) photon mapping host loop — BASELINE.json config 5. The reference has no such mode
(only the cutoff radius of its candidate list); this is a new capability built from the same
device stages: per pass `pm_emit_photons` (counter RNG, a fresh seed per pass) → `pm_map_build` →
`pm_gather_radius` at fixed hit points, followed by the radius / flux update of Hachisuka et al. 2008:

    N' = N + alpha * M          r'^2 = r^2 * N' / (N + M)          tau' = (tau + flux) * N' / (N + M)

with M photons found inside the current radius. Hit points are the first diffuse hits of the pixel-centre camera
rays. Multi-GPU: hit points replicated, each rank traces a disjoint share of every pass's photons
(first_photon = rank, photon_stride = world) and builds a map of its own shard; per pass one all-reduce (sum) of
(flux.rgb, count) — photons are partitioned, so the sums equal the single-GPU sums up to float addition order.
"""
from __future__ import annotations

import numpy as np


def camera_rays(pkg, camera, width: int, height: int) -> np.ndarray:
    """Pixel-centre rays of simpleRayGen's camera model (ray-tracer/cuda/deviceCode.cu:204-213), no jitter."""
    cam = np.asarray(camera).reshape(())
    px, py = np.meshgrid(np.arange(width, dtype=np.float32), np.arange(height, dtype=np.float32))
    su = ((px + np.float32(0.5)) / np.float32(width)).reshape(-1, 1)
    sv = ((py + np.float32(0.5)) / np.float32(height)).reshape(-1, 1)
    d = cam["dir_00"][None, :] + su * cam["dir_du"][None, :] + sv * cam["dir_dv"][None, :]
    d = (d / np.sqrt((d * d).sum(1, keepdims=True))).astype(np.float32)
    rays = np.zeros(width * height, pkg.RAY)
    rays["org"] = cam["pos"]
    rays["dir"] = d
    rays["tmin"], rays["tmax"] = 1e-3, 1e10
    return rays


def hit_points(pkg, scene, tri_mesh: np.ndarray, mats: np.ndarray, rays: np.ndarray):
    """(pos[n,3], brdf[n], albedo[n,3]): brdf = diffuse / pi at diffuse hits, 0 at misses / non-diffuse hits."""
    hits = scene.trace_closest(rays)
    ok = hits["prim"] >= 0
    mesh = np.where(ok, tri_mesh[np.maximum(hits["prim"], 0)], 0)
    m = mats[mesh]
    pos = (rays["org"] + rays["dir"] * np.where(ok, hits["t"], 0)[:, None].astype(np.float32)).astype(np.float32)
    brdf = np.where(ok, m["diffuse"] / np.float32(np.pi), 0).astype(np.float32)
    return np.ascontiguousarray(pos), brdf, np.ascontiguousarray(m["albedo"]).astype(np.float32)


def ppm_update(N, r2, tau, flux, count, alpha: float):
    """One radius / flux update, float32 arithmetic, in place. Pixels with count == 0 are unchanged."""
    M = count.astype(np.float32)
    hit = count > 0
    n_new = (N + np.float32(alpha) * M).astype(np.float32)
    ratio = np.where(hit, n_new / np.maximum(N + M, np.float32(1e-30)), np.float32(1)).astype(np.float32)
    r2 *= ratio
    tau += flux
    tau *= ratio[:, None]
    N[:] = np.where(hit, n_new, N)


def progressive_render(pkg, ctx, scene, tri_mesh, mats, lights, camera, width: int, height: int, passes: int,
                       photons_per_pass: int, r0: float, alpha: float = 0.7, max_depth: int = 8, seed: int = 1, group=None,
                       device=None):
    """Returns dict(radiance[H,W,3] float32, r2, N, tau, emitted). `group`: torch.distributed group for sharding."""
    rank, world = 0, 1
    if group is not None or _dist_ready():
        import torch.distributed as dist

        rank, world = dist.get_rank(group), dist.get_world_size(group)
    rays = camera_rays(pkg, camera, width, height)
    pos, brdf, albedo = hit_points(pkg, scene, np.asarray(tri_mesh), np.asarray(mats), rays)
    npx = width * height
    N = np.zeros(npx, np.float32)
    r2 = np.full(npx, np.float32(r0) * np.float32(r0), np.float32)
    tau = np.zeros((npx, 3), np.float32)
    emitted = 0
    for p in range(passes):
        g40, n, paths, _ = scene.emit_photons(lights, photons_per_pass, max_depth, False, seed=seed + p,
                                              rng_mode=pkg.RNG_COUNTER, first_photon=rank, photon_stride=world,
                                              capacity=(photons_per_pass // world + len(lights) + 1) * max(1, max_depth))
        emitted += photons_per_pass
        g44, _ = pkg.photons_to_maps(ctx, g40, np.zeros(0, pkg.PHOTON40))
        pmap = pkg.PhotonMap.build(ctx, g44) if len(g44) else None
        if pmap is not None:
            flux, count = pmap.gather_radius(pos, brdf, r2)
            pmap.close()
        else:
            flux, count = np.zeros((npx, 3), np.float32), np.zeros(npx, np.int32)
        if world > 1:
            flux, count = _allreduce(flux, count, group, device)
        ppm_update(N, r2, tau, flux, count, alpha)
    radiance = (tau / (np.float32(np.pi) * r2[:, None] * np.float32(max(emitted, 1)))).astype(np.float32) * albedo
    return dict(radiance=radiance.reshape(height, width, 3), r2=r2, N=N, tau=tau, emitted=emitted)


def _dist_ready() -> bool:
    try:
        import torch.distributed as dist

        return dist.is_available() and dist.is_initialized()
    except Exception:
        return False


def _allreduce(flux: np.ndarray, count: np.ndarray, group, device):
    """Per-pass exchange of config 5: sum of (flux.rgb, count) over ranks (4 values per pixel)."""
    import torch
    import torch.distributed as dist

    t = torch.from_numpy(np.concatenate([flux.astype(np.float64), count[:, None].astype(np.float64)], axis=1))
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t, group=group)
    t = t.cpu().numpy()
    return t[:, :3].astype(np.float32), np.rint(t[:, 3]).astype(np.int32)

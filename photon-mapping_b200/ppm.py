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


def paths_launched(lights, casted: int) -> int:
    """Paths one emission call launches over ALL ranks: per light int(power * int(casted / sum(power)))
    (photon-mapping/src/hostCode.cu:86,102-110; pm_emit.cu) — at most `casted`, less when the per-watt truncation
    bites. The radiance estimate is normalised by this number, not by the requested count."""
    total = float(np.sum(lights["power"])) if len(lights) else 0.0
    per_watt = int(int(casted) / total) if total > 0 else 0
    return sum(int(float(l["power"]) * per_watt) for l in lights)


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
    order = pkg.HitOrder(ctx, pos, brdf)  # the hit points never move: their spatial order is made once
    for p in range(passes):
        g40, n, paths, _ = scene.emit_photons(lights, photons_per_pass, max_depth, False, seed=seed + p,
                                              rng_mode=pkg.RNG_COUNTER, first_photon=rank, photon_stride=world,
                                              capacity=(photons_per_pass // world + len(lights) + 1) * max(1, max_depth))
        emitted += paths_launched(lights, photons_per_pass)
        g44, _ = pkg.photons_to_maps(ctx, g40, np.zeros(0, pkg.PHOTON40))
        pmap = pkg.PhotonMap.build(ctx, g44) if len(g44) else None
        if pmap is not None:
            flux, count = pmap.gather_radius(pos, brdf, r2, order=order)
            pmap.close()
        else:
            flux, count = np.zeros((npx, 3), np.float32), np.zeros(npx, np.int32)
        if world > 1:
            flux, count = _allreduce(flux, count, group, device)
        ppm_update(N, r2, tau, flux, count, alpha)
    order.close()
    radiance = (tau / (np.float32(np.pi) * r2[:, None] * np.float32(max(emitted, 1)))).astype(np.float32) * albedo
    return dict(radiance=radiance.reshape(height, width, 3), r2=r2, N=N, tau=tau, emitted=emitted)


def progressive_render_device(pkg, ctx, scene, tri_mesh, mats, lights, camera, width: int, height: int, passes: int,
                              photons_per_pass: int, r0: float, alpha: float = 0.7, max_depth: int = 8, seed: int = 1,
                              group=None, stats: list | None = None, cells: bool = False, cell_fraction: float = 0.25):
    """`progressive_render` with everything resident in HBM: hit points, per-pixel state, the pass's photons and its
    map never leave the device (the host loop above moves every pass's photons through numpy, ~80 B per photon over
    PCIe). Same arithmetic: the emission, record conversion, build and `pm_gather_radius` calls are the same C-ABI
    entry points on device pointers; the update is `ppm_update` in torch float32 (same IEEE operations element by
    element), so N, r² and tau equal the host loop's bit for bit on one GPU. Multi-GPU: photons of a pass sharded
    by launch index, one NCCL all-reduce (sum) of (flux.rgb, count) per pass — 16 B per pixel.
    `stats`: optional list that receives one dict of per-stage milliseconds per pass.
    `cells`: build a cells-only map per pass (pm_map_build_cells, cell = `cell_fraction` of the 90th-percentile radius of the lit pixels:
    blocks of at most 9 cells a side for nine hit points in ten)
    instead of the kd-tree map — the fixed-radius gather never touches the tree. Counts, hence N and r², are the
    same integers either way; flux sums differ in float addition order (photons are visited in another order)."""
    import ctypes
    import time

    import torch

    dev = torch.device("cuda", torch.cuda.current_device())
    rank, world = 0, 1
    if group is not None or _dist_ready():
        import torch.distributed as dist

        rank, world = dist.get_rank(group), dist.get_world_size(group)
    L = pkg.lib()
    lights = np.ascontiguousarray(lights, pkg.LIGHT)
    rays = camera_rays(pkg, camera, width, height)
    pos_h, brdf_h, albedo_h = hit_points(pkg, scene, np.asarray(tri_mesh), np.asarray(mats), rays)
    del rays
    npx = width * height
    pos, brdf = torch.from_numpy(pos_h).to(dev), torch.from_numpy(brdf_h).to(dev)
    N = torch.zeros(npx, dtype=torch.float32, device=dev)
    r2 = torch.full((npx,), float(np.float32(r0) * np.float32(r0)), dtype=torch.float32, device=dev)
    tau = torch.zeros((npx, 3), dtype=torch.float32, device=dev)
    fc = torch.zeros((npx, 4), dtype=torch.float32, device=dev)  # flux.rgb + count as float (exact below 2^24)
    flux = torch.empty((npx, 3), dtype=torch.float32, device=dev)
    count = torch.empty(npx, dtype=torch.int32, device=dev)
    cap = (photons_per_pass // world + len(lights) + 1) * max(1, max_depth)
    d40 = torch.empty((cap, 40), dtype=torch.uint8, device=dev)
    g44 = torch.empty((cap, 44), dtype=torch.uint8, device=dev)
    a32 = torch.tensor(alpha, dtype=torch.float32, device=dev)
    tiny, one = torch.tensor(1e-30, dtype=torch.float32, device=dev), torch.tensor(1.0, dtype=torch.float32, device=dev)
    emitted = 0
    order = pkg.HitOrder(ctx, pos, brdf)  # the hit points never move: their spatial order is made once
    torch.cuda.synchronize()
    for p in range(passes):
        t = [time.perf_counter()]

        def lap():
            torch.cuda.synchronize()
            t.append(time.perf_counter())

        cnt, npaths, nseg = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int64(0)
        pkg._check(L.pm_emit_photons(ctx.handle, scene.handle, lights.ctypes.data, len(lights), photons_per_pass, max_depth, 0,
                                     seed + p, pkg.RNG_COUNTER, rank, world, d40.data_ptr(), cap, ctypes.byref(cnt),
                                     ctypes.byref(npaths), ctypes.byref(nseg)))
        n = cnt.value
        emitted += paths_launched(lights, photons_per_pass)
        lap()
        if n > 0:
            pkg._check(L.pm_photons_to_maps(ctx.handle, d40.data_ptr(), n, None, 0, g44.data_ptr(), None))
            if cells:
                lit = N > 0
                r_med = float(torch.sqrt(torch.quantile(r2[lit][::61], 0.9) if bool(lit.any()) else r2[0]).item())
                pmap = pkg.PhotonMap.build_cells(ctx, g44[:n], cell_fraction * r_med)
            else:
                pmap = pkg.PhotonMap.build(ctx, g44[:n])
            lap()
            pmap.gather_radius(pos, brdf, r2, out_flux=flux, out_count=count, order=order)
            gk_ms, fallbacks = ctx.last_kernel_ms, ctx.last_fallback_count
            pmap.close()
        else:
            lap()
            flux.zero_()
            count.zero_()
            gk_ms, fallbacks = 0.0, 0
        lap()
        if world > 1:
            import torch.distributed as dist

            fc[:, :3] = flux
            fc[:, 3] = count.to(torch.float32)
            dist.all_reduce(fc, group=group)
            f_use, M = fc[:, :3], fc[:, 3]
            hit = M > 0
        else:
            f_use, M = flux, count.to(torch.float32)
            hit = count > 0
        # ppm_update, operation for operation
        n_new = N + a32 * M
        ratio = torch.where(hit, n_new / torch.maximum(N + M, tiny), one)
        r2 *= ratio
        tau += f_use
        tau *= ratio[:, None]
        N = torch.where(hit, n_new, N)
        lap()
        if stats is not None:
            stats.append({"pass": p, "stored": n, "paths": npaths.value, "segments": nseg.value, "emit_ms": 1e3 * (t[1] - t[0]),
                          "convert_build_ms": 1e3 * (t[2] - t[1]), "gather_ms": 1e3 * (t[3] - t[2]), "gather_kernels_ms": gk_ms,
                          "kd_range_fallbacks": fallbacks, "reduce_update_ms": 1e3 * (t[4] - t[3]),
                          "pass_ms": 1e3 * (t[4] - t[0])})
    N_h, r2_h, tau_h = N.cpu().numpy(), r2.cpu().numpy(), tau.cpu().numpy()
    radiance = (tau_h / (np.float32(np.pi) * r2_h[:, None] * np.float32(max(emitted, 1)))).astype(np.float32) * albedo_h
    return dict(radiance=radiance.reshape(height, width, 3), r2=r2_h, N=N_h, tau=tau_h, emitted=emitted)


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

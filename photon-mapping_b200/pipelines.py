"""Strong-scaling pipelines of the hot path over G ranks (one process per GPU), driven through the C ABI only — the
stages bench.py reports beside the C2 headline so that the driver measures the split SURVEY §8(e) names:

  C4-like   procedural 1 M-triangle room, a FIXED total of cast photons: emission sharded by launch index
            (pm_emit_photons_counted) -> ONE exchange (pm_exchange_photons: NCCL all-gather of count bytes and records,
            sort-free placement into the single-GPU order) -> sharded kd-tree build (pm_map_build_sharded) ->
            tile-sharded render (pm_render on this rank's pixel range) -> framebuffer merge (one all-reduce).
  C5-like   progressive fixed-radius passes at 3840x2160 on the Cornell box, a FIXED number of passes and photons per
            pass: per pass sharded emission -> cells-only map over the shard -> pm_gather_radius over the hit points ->
            pm_ppm_update (one all-reduce of 16 B per hit point + the radius / flux update in one kernel).

Every phase time is host-clocked with the device idle on both sides and the ranks aligned (barrier), max over ranks.
torch is used for device buffers and, under torchrun, for the barrier / max-reduction plumbing only."""
from __future__ import annotations

import ctypes
import os
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load_scene_npz(name):
    z = np.load(os.path.join(ROOT, "tests", "golden", f"scene_{name}.npz"))
    return z


class _Clock:
    """phase timer: barrier + device sync on both sides, max over ranks"""

    def __init__(self, world, dev):
        import torch

        self.world, self.dev, self.torch = world, dev, torch

    def _align(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist

            dist.barrier()
            self.torch.cuda.synchronize()

    def max_over_ranks(self, v: float) -> float:
        if self.world == 1:
            return float(v)
        import torch.distributed as dist

        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, f):
        self._align()
        t0 = time.perf_counter()
        r = f()
        self.torch.cuda.synchronize()
        return r, self.max_over_ranks(1e3 * (time.perf_counter() - t0))


def run_c4(pkg, ctx, comm, rank, world, casted_total=64_000_000, W=1920, H=1080, depth=8, iters=2, tris_u=500, tris_v=1000):
    import torch

    import __graft_entry__ as graft

    L = pkg.lib()
    dev = torch.device("cuda", ctx.device)
    clk = _Clock(world, dev)
    sc = graft.load_package_module("scenes").c4_room(tris_u, tris_v)
    lights = np.ascontiguousarray(sc["lights"], pkg.LIGHT)
    scene, scene_ms = clk.timed(lambda: pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"]))
    cam = pkg.make_camera(*sc["camera"], W, H)
    cam_c = np.ascontiguousarray(cam, pkg.CAMERA)
    sky = np.ones(3, np.float32)
    mg = graft.load_package_module("multigpu")
    sizes = mg.light_launch_sizes(lights, casted_total)
    my_paths = mg.rank_path_count(sizes, rank, world)
    cap_local = int(my_paths * 4.2) + 1024          # measured: 3.2 stored photons per path in this room
    cap_total = int(sum(sizes) * 4.2) + 1024
    shard = torch.empty((cap_local, 40), dtype=torch.uint8, device=dev)
    pcnt = torch.empty(my_paths + 16, dtype=torch.uint8, device=dev)
    g44 = torch.empty((cap_total, 44), dtype=torch.uint8, device=dev)
    fb = torch.zeros(W * H, dtype=torch.int32, device=dev)
    pb, pe = mg.shard_range(W * H, rank, world)
    best = None
    for it in range(iters + 1):  # the first iteration grows the context's scratch (cold); best of the rest
        cnt, npaths, nseg, total, coll_ms = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_float(0)

        def emit():
            pkg._check(L.pm_emit_photons_counted(ctx.handle, scene.handle, lights.ctypes.data, None, len(lights), casted_total, depth, 0, 1,
                                                 pkg.RNG_COUNTER, rank, world, shard.data_ptr(), pcnt.data_ptr(), cap_local,
                                                 ctypes.byref(cnt), ctypes.byref(npaths), ctypes.byref(nseg)))

        _, emit_ms = clk.timed(emit)
        walk_ms = clk.max_over_ranks(ctx.last_kernel_ms)

        def exchange():
            pkg._check(L.pm_exchange_photons(ctx.handle, comm.handle if comm else None, lights.ctypes.data, len(lights), casted_total,
                                             shard.data_ptr(), cnt.value, pcnt.data_ptr(), 1.0, g44.data_ptr(), cap_total,
                                             ctypes.byref(total), ctypes.byref(coll_ms)))

        _, ex_ms = clk.timed(exchange)
        n = total.value
        pmap, build_ms = clk.timed(lambda: pkg.PhotonMap.build_sharded_comm(ctx, comm, g44[:n]))

        def render():
            fb.zero_()
            ctx.after_torch(fb)
            pkg._check(L.pm_render(ctx.handle, scene.handle, lights.ctypes.data, len(lights), pmap.handle, None, cam_c.ctypes.data,
                                   W, H, 1, 1, sky.ctypes.data, pb, pe, fb.data_ptr()))

        _, render_ms = clk.timed(render)
        _, merge_ms = clk.timed(lambda: comm.allreduce_sum(fb) if comm else None)
        tree_sum = int(g44[:n].view(torch.int32).sum(dtype=torch.int64).item())
        lit = float((fb.view(H, W)[1:] != 0).float().mean().item())
        pmap.close()
        coll = clk.max_over_ranks(coll_ms.value)
        total_ms = emit_ms + ex_ms + build_ms + render_ms + merge_ms
        recv_bytes = (world - 1) * max(1, n // world) * 40  # records each rank receives from its peers (about)
        res = {"emit_ms": emit_ms, "walk_kernel_ms": walk_ms, "exchange_ms": ex_ms, "exchange_collective_ms": coll,
               "exchange_GBps_received_per_rank": (recv_bytes / (coll * 1e-3) / 1e9) if world > 1 and coll > 0 else None,
               "build_ms": build_ms, "render_ms": render_ms, "merge_ms": merge_ms, "total_ms": total_ms,
               "stored_photons": n, "paths_this_rank": npaths.value, "tree_checksum": tree_sum, "lit_fraction": lit}
        if it > 0 and (best is None or total_ms < best["total_ms"]):
            best = res
    # every rank must hold the same tree
    same = True
    if world > 1:
        import torch.distributed as dist

        t = torch.tensor([best["tree_checksum"]], dtype=torch.int64, device=dev)
        all_t = torch.empty(world, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(all_t, t)
        same = bool((all_t == all_t[0]).all().item())
    best.update({"workload": f"C4-like: {len(sc['tris'])} triangles, {casted_total} cast photons TOTAL (strong scaling), max_depth {depth}, "
                             f"counter RNG, {W}x{H} spp 1 depth 1; emit sharded by launch index -> pm_exchange_photons -> "
                             "pm_map_build_sharded -> pm_render per pixel range -> framebuffer all-reduce",
                 "ranks": world, "scene_create_ms": scene_ms, "cast_photons_per_s": casted_total / (best["total_ms"] * 1e-3),
                 "trees_identical_on_all_ranks": same, "scaling": "strong"})
    del shard, pcnt, g44, fb
    scene.close() if hasattr(scene, "close") else None
    return best


def run_c5(pkg, ctx, comm, rank, world, W=3840, H=2160, passes=8, per_pass=15_625_000, r0=0.5, alpha=0.7, depth=8):
    import torch

    import __graft_entry__ as graft

    L = pkg.lib()
    dev = torch.device("cuda", ctx.device)
    clk = _Clock(world, dev)
    ppm = graft.load_package_module("ppm")
    z = _load_scene_npz("cornell_box")
    mats = z["mats"].view(pkg.MATERIAL).reshape(-1)
    lights = np.ascontiguousarray(z["lights"].view(pkg.LIGHT).reshape(-1), pkg.LIGHT)
    scene = pkg.Scene(ctx, z["verts"], z["tris"], z["tri_mesh"], mats)
    cam = pkg.make_camera((80.0, 30.0, 0.0), (10.0, 20.0, 0.0), (0.0, 1.0, 0.0), 0.87, W, H)
    rays = ppm.camera_rays(pkg, cam, W, H)
    pos_h, brdf_h, _ = ppm.hit_points(pkg, scene, np.asarray(z["tri_mesh"]), mats, rays)
    del rays
    npx = W * H
    pos, brdf = torch.from_numpy(pos_h).to(dev), torch.from_numpy(brdf_h).to(dev)
    N = torch.zeros(npx, dtype=torch.float32, device=dev)
    r2 = torch.full((npx,), float(np.float32(r0) * np.float32(r0)), dtype=torch.float32, device=dev)
    tau = torch.zeros((npx, 3), dtype=torch.float32, device=dev)
    flux = torch.empty((npx, 3), dtype=torch.float32, device=dev)
    count = torch.empty(npx, dtype=torch.int32, device=dev)
    cap = (per_pass // world + len(lights) + 1) * depth
    d40 = torch.empty((cap, 40), dtype=torch.uint8, device=dev)
    g44 = torch.empty((cap, 44), dtype=torch.uint8, device=dev)
    order = pkg.HitOrder(ctx, pos, brdf)  # the hit points never move: pm_hit_order_create once, reused by every pass
    sample = slice(0, npx, 61)  # the cell size is a speed knob only: a strided sample of the radii is enough
    stats = []
    t_all0 = None
    for p in range(passes + 2):  # two warm-up passes (scratch and pool growth) are run and not counted
        if p == 2:
            torch.cuda.synchronize()
            clk._align()
            t_all0 = time.perf_counter()
        cnt, npaths, nseg, coll_ms = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_float(0)
        t = [time.perf_counter()]

        def lap():
            torch.cuda.synchronize()
            t.append(time.perf_counter())

        pkg._check(L.pm_emit_photons(ctx.handle, scene.handle, lights.ctypes.data, len(lights), per_pass, depth, 0, 1 + p,
                                     pkg.RNG_COUNTER, rank, world, d40.data_ptr(), cap, ctypes.byref(cnt), ctypes.byref(npaths),
                                     ctypes.byref(nseg)))
        n = cnt.value
        lap()
        lit = N[sample] > 0
        # cell = a quarter of the 90th-percentile radius: blocks of at most 9 cells a side for nine hit points in ten
        r_med = float(torch.sqrt(torch.quantile(r2[sample][lit], 0.9) if bool(lit.any()) else r2[0]).item())
        pkg._check(L.pm_photons_to_maps(ctx.handle, d40.data_ptr(), n, None, 0, g44.data_ptr(), None))
        pmap = pkg.PhotonMap.build_cells(ctx, g44[:n], 0.25 * r_med)
        lap()
        pmap.gather_radius(pos, brdf, r2, out_flux=flux, out_count=count, order=order)
        gk = ctx.last_kernel_ms
        pmap.close()
        lap()
        pkg._check(L.pm_ppm_update(ctx.handle, comm.handle if comm else None, npx, flux.data_ptr(), count.data_ptr(), alpha,
                                   N.data_ptr(), r2.data_ptr(), tau.data_ptr(), ctypes.byref(coll_ms)))
        lap()
        if p >= 2:
            stats.append({"emit_ms": 1e3 * (t[1] - t[0]), "map_ms": 1e3 * (t[2] - t[1]), "gather_ms": 1e3 * (t[3] - t[2]),
                          "gather_kernels_ms": gk, "update_ms": 1e3 * (t[4] - t[3]), "allreduce_ms": coll_ms.value, "stored": n})
    torch.cuda.synchronize()
    loop_ms = clk.max_over_ranks(1e3 * (time.perf_counter() - t_all0))
    order.close()
    mean = lambda k: clk.max_over_ranks(float(np.mean([s[k] for s in stats])))
    lit_px = int((N > 0).sum().item())
    res = {"workload": f"C5-like: Cornell box, {W}x{H} hit points, {passes} passes x {per_pass} cast photons (fixed total: strong "
                       f"scaling), r0 {r0}, alpha {alpha}, max_depth {depth}; per pass: sharded emission -> cells-only map of the shard "
                       "-> pm_gather_radius_ordered over all hit points (one pm_hit_order for the run) -> pm_ppm_update (all-reduce of 16 B per hit point + update kernel)",
           "ranks": world, "passes": passes, "loop_ms": loop_ms, "pass_ms": loop_ms / passes,
           "cast_photons_per_s": passes * per_pass / (loop_ms * 1e-3), "emit_ms": mean("emit_ms"), "map_ms": mean("map_ms"),
           "gather_ms": mean("gather_ms"), "gather_kernels_ms": mean("gather_kernels_ms"), "update_ms": mean("update_ms"),
           "allreduce_ms": mean("allreduce_ms"),
           "allreduce_busbw_GBps": (2 * (world - 1) / world * npx * 16 / (mean("allreduce_ms") * 1e-3) / 1e9) if world > 1 else None,
           "stored_photons_per_pass_this_rank": float(np.mean([s["stored"] for s in stats])), "lit_hit_points": lit_px,
           "state_checksum": float(N.double().sum().item()), "scaling": "strong"}
    return res


def run(pkg, ctx, rank: int, local_rank: int, world: int, c4_total: int = 64_000_000):
    """Both pipelines; returns the dict for the bench line on rank 0 (None elsewhere)."""
    comm = pkg.Comm.from_torch_distributed(ctx) if world > 1 else None
    out = {}
    for name, fn in (("c4_pipeline", lambda: run_c4(pkg, ctx, comm, rank, world, casted_total=c4_total)),
                     ("c5_progressive", lambda: run_c5(pkg, ctx, comm, rank, world))):
        try:
            out[name] = fn()
        except Exception as e:  # a failing stage must not cost the headline line; the other ranks fail the same way
            out[name] = {"error": repr(e)}
        import torch

        torch.cuda.empty_cache()
    if comm:
        comm.close()
    return out if rank == 0 else None

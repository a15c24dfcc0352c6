#!/usr/bin/env python
"""C5 (BASELINE.json configs[4]): fixed-radius progressive photon mapping, Cornell box at 3840x2160, 64 passes x
15 625 000 cast photons = 1 G photons, r0 = 0.5, alpha = 0.7, everything resident in HBM
(ppm.progressive_render_device). One process per GPU: every pass's photons are sharded by launch index, hit points
are replicated, one NCCL all-reduce of (flux.rgb, count) per pass.

  python tools/c5_run.py [passes] [photons_per_pass] [W] [H] [r0] [cells|tree]
  torchrun --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29512 tools/c5_run.py ...

No reference implementation exists; the parity tests are tests/test_ppm.py. Checked here:
radii never grow, N grows by alpha * M, radiance finite; with G > 1 a small case is compared with the same loop run
unsharded on every rank (N and r² exact — counts are integers —, tau to float-addition order).
Rank 0 writes gpurun_out/c5_run_<G>gpu.json; pass times are max over ranks."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
passes = int(sys.argv[1]) if len(sys.argv) > 1 else 64
per_pass = int(sys.argv[2]) if len(sys.argv) > 2 else 15_625_000
W = int(sys.argv[3]) if len(sys.argv) > 3 else 3840
H = int(sys.argv[4]) if len(sys.argv) > 4 else 2160
r0 = float(sys.argv[5]) if len(sys.argv) > 5 else 0.5
cells = (sys.argv[6] != "tree") if len(sys.argv) > 6 else True  # per-pass map: cells-only (default) or the kd-tree map
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
group = None
if world > 1:
    import torch.distributed as dist

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29512")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
pkg = graft.load_package()
ppm = graft.load_package_module("ppm")
ctx = pkg.Context(local)
z = np.load(os.path.join(ROOT, "tests", "golden", "scene_cornell_box.npz"))
mats = z["mats"].view(pkg.MATERIAL).reshape(-1)
scene = pkg.Scene(ctx, z["verts"], z["tris"], z["tri_mesh"], mats)
lights = z["lights"].view(pkg.LIGHT).reshape(-1)
alpha, depth = 0.7, 8
res = {"ranks": world, "fb": [W, H], "passes": passes, "photons_per_pass": per_pass, "r0": r0, "alpha": alpha, "max_depth": depth,
       "per_pass_map": "cells-only (pm_map_build_cells)" if cells else "kd-tree (pm_map_build)"}


def run(w, h, n_pass, n_photons, radius, sharded, stats=None):
    cam = pkg.make_camera((80.0, 30.0, 0.0), (10.0, 20.0, 0.0), (0.0, 1.0, 0.0), 0.87, w, h)
    if sharded or world == 1:
        return ppm.progressive_render_device(pkg, ctx, scene, z["tri_mesh"], mats, lights, cam, w, h, n_pass, n_photons, radius, alpha,
                                             depth, seed=1, stats=stats, cells=cells, cell_fraction=float(os.environ.get('PM_PPM_CELL_FRAC', '0.25')))
    # unsharded on this rank although a process group exists: a one-rank group of our own
    return ppm.progressive_render_device(pkg, ctx, scene, z["tri_mesh"], mats, lights, cam, w, h, n_pass, n_photons, radius, alpha,
                                         depth, seed=1, stats=stats, group=solo, cells=cells, cell_fraction=float(os.environ.get('PM_PPM_CELL_FRAC', '0.25')))


if world > 1:
    solos = [dist.new_group([r]) for r in range(world)]
    solo = solos[rank]
    a = run(160, 90, 3, 200_000, 2.0, True)
    b = run(160, 90, 3, 200_000, 2.0, False)
    ok = np.array_equal(a["N"], b["N"]) and np.array_equal(a["r2"], b["r2"]) and np.allclose(a["tau"], b["tau"], rtol=2e-4, atol=1e-6)
    flag = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    res["small_case_sharded_equals_unsharded"] = bool(flag.item())
    assert flag.item() == 1

stats = []
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.perf_counter()
out = run(W, H, passes, per_pass, r0, True, stats)
torch.cuda.synchronize()
total_s = time.perf_counter() - t0
keys = ["emit_ms", "convert_build_ms", "gather_ms", "gather_kernels_ms", "reduce_update_ms", "pass_ms"]
per = np.array([[s[k] for k in keys] for s in stats], np.float64)
stored = np.array([s["stored"] for s in stats], np.float64)
fallbacks = np.array([s["kd_range_fallbacks"] for s in stats], np.float64)
if world > 1:
    tt = torch.from_numpy(per).to(dev)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    per = tt.cpu().numpy()
    ts = torch.from_numpy(stored).to(dev)
    dist.all_reduce(ts)
    stored = ts.cpu().numpy()
    tw = torch.tensor([total_s], dtype=torch.float64, device=dev)
    dist.all_reduce(tw, op=dist.ReduceOp.MAX)
    total_s = float(tw.item())
r2 = out["r2"]
lit = out["N"] > 0
assert np.isfinite(out["radiance"]).all() and (r2 <= np.float32(r0) ** 2).all() and lit.any()
res.update({
    "cast_photons_total": passes * per_pass, "stored_photons_total": float(stored.sum()),
    "loop_s_including_hit_points": total_s, "passes_s": float(per[:, 5].sum()) / 1e3,
    "cast_photons_per_s": passes * per_pass / (per[:, 5].sum() / 1e3),
    "mean_ms": {k: float(per[:, i].mean()) for i, k in enumerate(keys)},
    "first_pass_ms": {k: float(per[0, i]) for i, k in enumerate(keys)},
    "last_pass_ms": {k: float(per[-1, i]) for i, k in enumerate(keys)},
    "kd_range_fallbacks_rank0_first_last": [float(fallbacks[0]), float(fallbacks[-1])],
    "pixels_with_photons": float(lit.mean()), "median_radius_last": float(np.sqrt(np.median(r2[lit]))),
    "mean_N_last": float(out["N"][lit].mean()), "radiance_mean": float(out["radiance"].mean()),
})
if rank == 0:
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", f"c5_run_{world}gpu{'' if cells else '_tree'}.json"), "w"), indent=1)
    print(json.dumps(res))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()

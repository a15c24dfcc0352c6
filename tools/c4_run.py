#!/usr/bin/env python
"""C4-like pass over G ranks (one process per GPU, NCCL): procedural 1 M-triangle scene replicated, stage 1 sharded by
launch index (rank r traces r, r + G, ...), ONE exchange (all_gather of counts, then of the padded 40-byte records and
their 64-bit keys; a sort by key restores the single-GPU order), replicated map build, pixel-sharded render.

  torchrun --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29511 tools/c4_run.py [casted_per_rank] [W] [H]

Checks (no oracle on the box: properties only):
  * small case (400 k cast photons): the merged array of the G ranks equals, byte for byte, the unsharded emission
    every rank runs itself; the merged framebuffer equals the unsharded render;
  * full case: every rank holds the same merged array (64-bit checksum all-gathered), keys strictly increase.
Times are max over ranks, device-synchronised on both sides. Rank 0 writes gpurun_out/c4_run_<G>gpu.json."""
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
per_rank = int(sys.argv[1]) if len(sys.argv) > 1 else 25_000_000
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
H = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29511")
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
pkg = graft.load_package()
mg = graft.load_package_module("multigpu")
ctx = pkg.Context(local)
L = pkg.lib()


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
    # wound so that the geometric normals point INTO the room: the reference scatters about the unflipped normal
    # (common/cuda/helpers.h:45-47,76-85), so a room whose walls face outwards loses every photon at its first wall
    tris = np.array([t for q in quads for t in ((q[0], q[2], q[1]), (q[0], q[3], q[2]))], np.int32)
    return c, tris


def sync_max(ms: float) -> float:
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed(f):
    """Host clock around a call with the device idle on both sides and the ranks aligned; max over ranks."""
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = f()
    torch.cuda.synchronize()
    return r, sync_max(1e3 * (time.perf_counter() - t0))


def emit_shard(scene, lights, casted, depth, first, stride, cap):
    """pm_emit_photons_keyed into device buffers: (records u8 [n, 40], keys i64 [n], paths, segments, walk kernel ms)."""
    rec = torch.empty((cap, 40), dtype=torch.uint8, device=dev)
    key = torch.empty(cap, dtype=torch.int64, device=dev)
    cnt, npaths, nseg = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int64(0)
    pkg._check(L.pm_emit_photons_keyed(ctx.handle, scene.handle, lights.ctypes.data, len(lights), casted, depth, 0, 1, pkg.RNG_COUNTER,
                                       first, stride, rec.data_ptr(), key.data_ptr(), cap, ctypes.byref(cnt), ctypes.byref(npaths),
                                       ctypes.byref(nseg)))
    return rec[: cnt.value], key[: cnt.value], npaths.value, nseg.value, ctx.last_kernel_ms


def exchange(rec, key):
    """multigpu.exchange_photons + its timing: (records, keys, counts, all_gather ms, padded capacity)."""
    tm = {}
    recs, keys, counts = mg.exchange_photons(rec, key, timing=tm)
    torch.cuda.synchronize()
    return recs, keys, counts, tm["all_gather_ms"], tm["padded_capacity"]


def checksum(t: torch.Tensor) -> int:
    v = t.reshape(-1).view(torch.int64)
    a = torch.zeros((), dtype=torch.int64, device=v.device)
    b = torch.zeros((), dtype=torch.int64, device=v.device)
    step = 1 << 26  # position-weighted sum in pieces: the maps of a full C4 run leave little free memory
    for lo in range(0, v.numel(), step):
        c = v[lo:lo + step]
        a += (c * torch.arange(lo + 1, lo + 1 + c.numel(), dtype=torch.int64, device=v.device)).sum()
        b += c.sum()
    return int((a ^ b).item())


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
lights = np.ascontiguousarray(lights, pkg.LIGHT)
scene, scene_ms = timed(lambda: pkg.Scene(ctx, verts, tris, tri_mesh, mats))
res = {"ranks": world, "triangles": int(len(tris)), "scene_create_ms": scene_ms, "fb": [W, H]}
cam = pkg.make_camera((0.0, 20.0, -19.0), (0.0, 14.0, 0.0), (0.0, 1.0, 0.0), 0.87, W, H)
depth = 8


def to_map(rec40):
    g44 = torch.empty((rec40.shape[0], 44), dtype=torch.uint8, device=dev)
    pkg._check(L.pm_photons_to_maps(ctx.handle, rec40.data_ptr(), rec40.shape[0], None, 0, g44.data_ptr(), None))
    return g44


# ---- small case: sharded == unsharded, byte for byte --------------------------------------------------
small = 400_000
cap_s = small * 5
full_rec, full_key, _, _, _ = emit_shard(scene, lights, small, depth, 0, 1, cap_s)
sh_rec, sh_key, _, _, _ = emit_shard(scene, lights, small, depth, rank, world, cap_s)
m_rec, m_key, counts_s, _, _ = exchange(sh_rec, sh_key)
ok = bool(torch.equal(m_rec, full_rec)) and bool(torch.equal(m_key, full_key))
g44 = to_map(full_rec)
gmap = pkg.PhotonMap.build(ctx, g44)
cam_s = pkg.make_camera((0.0, 20.0, -19.0), (0.0, 14.0, 0.0), (0.0, 1.0, 0.0), 0.87, 256, 144)
fb_full = scene.render(lights, gmap, None, cam_s, 256, 144, 1, 1)
fb_merged = mg.render_sharded(pkg, scene, lights, gmap, None, cam_s, 256, 144, 1, 1, device=dev)
ok_fb = bool(np.array_equal(fb_full, fb_merged))
flag = torch.tensor([int(ok), int(ok_fb)], dtype=torch.int64, device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
res["small_case"] = {"casted": small, "stored": int(full_rec.shape[0]), "per_rank_counts": counts_s,
                     "merged_equals_single_gpu_bytes": bool(flag[0].item()), "merged_fb_equals_single_gpu": bool(flag[1].item())}
assert flag.min().item() == 1, res["small_case"]
del full_rec, full_key, sh_rec, sh_key, m_rec, m_key, gmap, g44

# ---- full case --------------------------------------------------------------------------------------
casted = per_rank * world
cap = per_rank * 5
for it in range(2):  # the first call grows the context's scratch
    (rec, key, paths, segs, walk_ms), emit_ms = timed(lambda: emit_shard(scene, lights, casted, depth, rank, world, cap))
    if it == 0:
        del rec, key
tot = torch.tensor([paths, segs, rec.shape[0]], dtype=torch.int64, device=dev)
dist.all_reduce(tot)
res["stage1"] = {"casted_total": casted, "paths_total": int(tot[0]), "segments_total": int(tot[1]), "stored_total": int(tot[2]),
                 "call_ms_max_over_ranks": emit_ms, "walk_kernel_ms_max": sync_max(walk_ms),
                 "paths_per_s_aggregate": int(tot[0]) / (emit_ms * 1e-3), "stored_photons_per_s_aggregate": int(tot[2]) / (emit_ms * 1e-3)}
(m_rec, m_key, counts, coll_ms, padcap), ex_ms = timed(lambda: exchange(rec, key))
del rec, key
coll_ms = sync_max(coll_ms)
coll_bytes = (world - 1) * padcap * 48  # what every rank receives from its peers (records + keys, padded)
sums = torch.tensor([checksum(m_rec), checksum(m_key)], dtype=torch.int64, device=dev)
all_sums = torch.empty((world, 2), dtype=torch.int64, device=dev)
dist.all_gather_into_tensor(all_sums, sums)
same = bool((all_sums == all_sums[0]).all().item())
incr = bool((m_key[1:] > m_key[:-1]).all().item())
res["exchange"] = {"per_rank_counts": counts, "all_gather_ms_max": coll_ms, "all_gather_bytes_received_per_rank": coll_bytes,
                   "all_gather_GBps_per_rank": coll_bytes / (coll_ms * 1e-3) / 1e9, "exchange_with_key_sort_ms_max": ex_ms,
                   "replicas_identical_checksum": same, "keys_strictly_increase": incr}
assert same and incr
del m_key
torch.cuda.empty_cache()
g44, conv_ms = timed(lambda: to_map(m_rec))
del m_rec
torch.cuda.empty_cache()  # the library allocates with cudaMalloc: hand torch's cached blocks back first
def build(t):
    """sharded build: levels below log2(world) split over the ranks, node tables summed with two NCCL all-reduces"""
    if os.environ.get("PM_C4_BUILD", "sharded") == "replicated" or world & (world - 1):
        return pkg.PhotonMap.build(ctx, t)
    return pkg.PhotonMap.build_sharded(ctx, t, rank, world, lambda x: dist.all_reduce(x))


b0 = None
if g44.shape[0] <= 350_000_000:  # warm build on a copy (grows the context's scratch); skipped when memory is the concern
    g_copy = g44.clone()
    gmap, b0 = timed(lambda: build(g_copy))
    del gmap, g_copy
gmap, build_ms = timed(lambda: build(g44))
tree_sum = torch.tensor([checksum(g44)], dtype=torch.int64, device=dev)
tree_sums = torch.empty(world, dtype=torch.int64, device=dev)
dist.all_gather_into_tensor(tree_sums, tree_sum)
res["stage2"] = {"photons": int(g44.shape[0]), "to_maps_ms": conv_ms, "build_ms_first": b0, "build_ms": build_ms,
                 "mode": os.environ.get("PM_C4_BUILD", "sharded"), "tree_checksum": int(tree_sum.item()),
                 "trees_identical_on_all_ranks": bool((tree_sums == tree_sums[0]).all().item())}
assert res["stage2"]["trees_identical_on_all_ranks"]
pb, pe = mg.shard_range(W * H, rank, world)
for it in range(2):
    _, own_ms = timed(lambda: scene.render(lights, gmap, None, cam, W, H, 1, 1, (1.0, 1.0, 1.0), pb, pe))
gk_ms, fallbacks = sync_max(ctx.last_kernel_ms), ctx.last_fallback_count
for it in range(2):
    fb, render_ms = timed(lambda: mg.render_sharded(pkg, scene, lights, gmap, None, cam, W, H, 1, 1, device=dev))
res["stage3_sharded"] = {"pixels_per_rank": pe - pb, "render_own_range_ms_max": own_ms, "gather_kernels_ms_last_wave_max": gk_ms,
                         "kd_fallbacks_rank0": fallbacks, "render_plus_merge_ms_max": render_ms,
                         "lit_fraction": float((fb[1:] != 0).mean()), "pixels_per_s_aggregate": W * H / (render_ms * 1e-3)}
res["total_ms"] = emit_ms + ex_ms + conv_ms + build_ms + render_ms
res["peak_mem_GB_rank0"] = torch.cuda.max_memory_allocated() / 1e9
if rank == 0:
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", f"c4_run_{world}gpu.json"), "w"), indent=1)
    print(json.dumps(res))
dist.barrier()
dist.destroy_process_group()

"""Multi-GPU host logic: one process per GPU, `torch.distributed` for the plumbing (NCCL over NVLink on the
GPU box, gloo in the CPU tests). The path shards without a data-path collective except for ONE exchange
between stage 1 and stage 2:

  stage 1  rank r traces launch indices r, r + G, r + 2G, ... of every light (pm_emit_photons_keyed with
           first_photon = r, photon_stride = G); scene and BVH are replicated;
  exchange all_gather of the per-rank counts, then all_gather of the padded 40-byte records and their
           64-bit keys; sorting by key restores exactly the single-GPU output order (light, launch index,
           bounce), so the replicated photon map is bit-identical to the 1-GPU map. (The library's own
           exchange, pm_exchange_photons in csrc/pm_comm.cu, needs neither keys nor a sort: one count byte
           per path and two prefix sums place every record; `place_by_counts` below states it in numpy.)
  stage 2  every rank builds the same map (replicated build);
  stage 3  pixels (or queries) shard by contiguous range, no communication; framebuffers are summed
           (disjoint ranges, zero elsewhere) or gathered on rank 0.
"""
from __future__ import annotations

import numpy as np

PHOTON40_BYTES = 40


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous share [begin, end) of `total` items for `rank` (first `total % world` ranks get one more)."""
    base, rem = divmod(int(total), int(world))
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def light_launch_sizes(lights: np.ndarray, casted: int) -> list[int]:
    """Launch size per light as the reference computes it (photon-mapping/src/hostCode.cu:86,102-110):
    photonsPerWatt = int(casted / sum(power)); per light int(power * photonsPerWatt)."""
    total = float(np.sum(lights["power"]))
    per_watt = int(int(casted) / total) if len(lights) and total > 0 else 0
    return [int(float(l["power"]) * per_watt) for l in lights]


def rank_path_count(sizes: list[int], rank: int, world: int) -> int:
    """Number of paths rank `rank` traces: launch indices rank, rank + world, ... below each light's size."""
    return sum((n - rank + world - 1) // world if n > rank else 0 for n in sizes)


def merge_shards(records: list[np.ndarray], keys: list[np.ndarray]) -> tuple[np.ndarray, np.ndarray]:
    """Concatenate per-rank (records, keys) and sort by key: the single-GPU order."""
    rec = np.concatenate(records) if records else np.zeros(0, np.uint8)
    key = np.concatenate(keys) if keys else np.zeros(0, np.uint64)
    order = np.argsort(key, kind="stable")
    return rec[order], key[order]


def exchange_layout(sizes: list[int], world: int):
    """The path layout of a sharded emission, as pm_exchange_photons lays it out (csrc/pm_comm.cu make_layout/locate):
    global path g = light_base[l] + i belongs to rank i % world and is that rank's local path lf[rank][l] + i // world.
    Returns (rank_of[g], local_of[g], paths_per_rank)."""
    rank_of, local_of = [], []
    lf = [0] * world
    for n in sizes:
        i = np.arange(int(n), dtype=np.int64)
        r = i % world
        rank_of.append(r)
        local_of.append(np.asarray(lf, np.int64)[r] + i // world)
        for k in range(world):
            lf[k] += (n - k + world - 1) // world if n > k else 0
    cat = lambda a: np.concatenate(a) if a else np.zeros(0, np.int64)
    return cat(rank_of), cat(local_of), lf


def place_by_counts(counts: list[np.ndarray], records: list[np.ndarray], sizes: list[int]) -> np.ndarray:
    """Sort-free merge of per-rank shards (numpy statement of pm_comm.cu exchange_place): `counts[r]` holds one byte per
    path rank r traced (photons stored on it), `records[r]` that rank's records in its own emission order. Two
    prefix sums (global path order, per-rank order) give every path's source and destination offsets; the result is
    the single-GPU record order without any key."""
    world = len(counts)
    rank_of, local_of, per_rank = exchange_layout(sizes, world)
    for r in range(world):
        assert len(counts[r]) == per_rank[r], "count bytes do not match the launch layout"
    c = np.array([counts[r][p] for r, p in zip(rank_of, local_of)], np.int64) if len(rank_of) else np.zeros(0, np.int64)
    goff = np.concatenate([[0], np.cumsum(c)])
    roff = [np.concatenate([[0], np.cumsum(counts[r].astype(np.int64))]) for r in range(world)]
    width = records[0].shape[1] if records and records[0].ndim == 2 else PHOTON40_BYTES
    out = np.zeros((int(goff[-1]), width), np.uint8)
    for g in np.nonzero(c)[0]:
        r, p = int(rank_of[g]), int(local_of[g])
        out[goff[g]:goff[g + 1]] = records[r][roff[r][p]:roff[r][p + 1]]
    return out


def allgather_photons(local_records, local_keys, group=None, device=None):
    """All-gather variable-length photon shards. `local_records`: numpy structured array (40-byte records) or
    a uint8 torch tensor of n*40 bytes; `local_keys`: uint64 numpy / int64 torch of n keys. Returns the merged
    (records uint8 [N, 40], keys int64 [N]) torch tensors in single-GPU order, replicated on every rank."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    if isinstance(local_records, np.ndarray):
        rec = torch.from_numpy(np.ascontiguousarray(local_records).view(np.uint8).reshape(-1, PHOTON40_BYTES).copy())
        key = torch.from_numpy(np.ascontiguousarray(local_keys).view(np.int64).copy())
    else:
        rec, key = local_records.reshape(-1, PHOTON40_BYTES), local_keys.reshape(-1)
    if device is not None:
        rec, key = rec.to(device), key.to(device)
    n = torch.tensor([rec.shape[0]], dtype=torch.int64, device=rec.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    cap = max(max(counts), 1)
    pad_rec = torch.zeros((cap, PHOTON40_BYTES), dtype=torch.uint8, device=rec.device)
    pad_key = torch.zeros(cap, dtype=torch.int64, device=rec.device)
    pad_rec[: rec.shape[0]] = rec
    pad_key[: key.shape[0]] = key
    all_rec = [torch.empty_like(pad_rec) for _ in range(world)]
    all_key = [torch.empty_like(pad_key) for _ in range(world)]
    dist.all_gather(all_rec, pad_rec, group=group)
    dist.all_gather(all_key, pad_key, group=group)
    rec = torch.cat([r[:c] for r, c in zip(all_rec, counts)])
    key = torch.cat([k[:c] for k, c in zip(all_key, counts)])
    # keys are < 2^63 (light < 128), so the signed order equals the unsigned one
    order = torch.sort(key, stable=True).indices
    return rec[order], key[order]


def exchange_photons(rec, key, group=None, timing: dict | None = None):
    """The path's one collective for tensors that stay where they are (HBM under NCCL, host memory under gloo): counts,
    then ONE all_gather_into_tensor each for the padded 40-byte records and the 64-bit keys, then a sort by key — the
    single-GPU order — with the records permuted as five 8-byte columns. `rec`: uint8 [n, 40], `key`: int64 [n].
    Returns (records uint8 [N, 40], keys int64 [N], per-rank counts); `timing` receives the collective's device
    milliseconds (CUDA tensors only) and the padded per-rank capacity."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    dev = rec.device
    rec, key = rec.reshape(-1, PHOTON40_BYTES), key.reshape(-1)
    n = torch.tensor([rec.shape[0]], dtype=torch.int64, device=dev)
    counts = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(counts, n, group=group)
    counts = counts.tolist()
    cap = max(max(counts), 1)
    # records travel as five int64 columns (40 bytes, 8-byte aligned): flat 1-D int64 buffers suit every backend
    pad_rec = torch.zeros(cap * 5, dtype=torch.int64, device=dev)
    pad_key = torch.zeros(cap, dtype=torch.int64, device=dev)
    pad_rec[: rec.shape[0] * 5] = rec.contiguous().view(torch.int64).reshape(-1)
    pad_key[: key.shape[0]] = key
    all_rec = torch.empty(world * cap * 5, dtype=torch.int64, device=dev)
    all_key = torch.empty(world * cap, dtype=torch.int64, device=dev)
    ev = None
    if dev.type == "cuda":
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record()
    dist.all_gather_into_tensor(all_rec, pad_rec, group=group)
    dist.all_gather_into_tensor(all_key, pad_key, group=group)
    if ev:
        ev[1].record()
    del pad_rec, pad_key
    all_rec, all_key = all_rec.view(world, cap, 5), all_key.view(world, cap)
    keys = torch.cat([all_key[r, :c] for r, c in enumerate(counts)])
    del all_key
    # keys are < 2^63 (light < 128), so the signed order equals the unsigned one
    order = torch.sort(keys, stable=True).indices
    keys = keys[order]
    recs = torch.cat([all_rec[r, :c] for r, c in enumerate(counts)])
    del all_rec
    recs = recs[order].view(torch.uint8).view(-1, PHOTON40_BYTES)
    if timing is not None:
        timing["padded_capacity"] = cap
        if ev:
            torch.cuda.synchronize()
            timing["all_gather_ms"] = ev[0].elapsed_time(ev[1])
    return recs, keys, counts


def emit_sharded(pkg, scene, lights, casted: int, max_depth: int, caustics: bool, seed: int = 0, rng_mode: int = 0,
                 group=None, device=None):
    """Stage 1 over all ranks + the exchange. Returns the full PHOTON40 numpy array (replicated, single-GPU order)."""
    import torch.distributed as dist

    rank, world = dist.get_rank(group), dist.get_world_size(group)
    rec, keys, _, _ = scene.emit_photons_keyed(lights, casted, max_depth, caustics, seed, rng_mode, rank, world)
    all_rec, _ = allgather_photons(rec, keys, group, device)
    return all_rec.cpu().numpy().reshape(-1).view(pkg.PHOTON40)


def render_sharded(pkg, scene, lights, gmap, cmap, camera, width: int, height: int, spp: int, depth: int, sky=(1.0, 1.0, 1.0),
                   group=None, device=None):
    """Stage 3 over all ranks: rank r renders its pixel range; the framebuffers (zero outside the range) are
    summed. Returns uint32[H, W] on every rank."""
    import torch
    import torch.distributed as dist

    rank, world = dist.get_rank(group), dist.get_world_size(group)
    b, e = shard_range(width * height, rank, world)
    fb = scene.render(lights, gmap, cmap, camera, width, height, spp, depth, sky, b, e)
    t = torch.from_numpy(fb.view(np.int32).astype(np.int64))  # disjoint ranges: the sum is a merge
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t, group=group)
    return (t.cpu().numpy() & 0xFFFFFFFF).astype(np.uint32).reshape(height, width)

"""Host-side multi-GPU logic on the CPU: world_size-2 gloo process group. The device work is
replaced by a deterministic fake emitter with the same contract as pm_emit_photons_keyed (records + keys
for launch indices first, first + stride, ...), so what is tested is the sharding arithmetic, the
variable-length all-gather and the key merge that restores the single-GPU order."""
import os
import socket

import numpy as np
import pytest

import __graft_entry__ as graft

LIGHT = np.dtype([("pos", "<f4", 3), ("rgb", "<f4", 3), ("power", "<f8")])


def _lights():
    l = np.zeros(3, LIGHT)
    l["power"] = [10.0, 2.5, 7.5]
    return l


def fake_emit(sizes, first, stride):
    """records (uint8[n,40]) and keys (uint64[n]) for launch indices first, first+stride, ... of every light."""
    rec, keys = [], []
    for l, n in enumerate(sizes):
        for i in range(first, n, stride):
            stores = (i * 2654435761 + l * 97) % 4  # 0..3 stored photons on this path
            for k in range(stores):
                key = (l << 56) | (i << 8) | k
                keys.append(key)
                rec.append(np.frombuffer(np.random.default_rng(key % (2**32)).bytes(40), np.uint8))
    if not rec:
        return np.zeros((0, 40), np.uint8), np.zeros(0, np.uint64)
    return np.stack(rec), np.array(keys, np.uint64)


def fake_counts(sizes, first, stride):
    """One count byte per path of the same fake emission, in this rank's emission order (pm_emit_photons_counted)."""
    return np.array([(i * 2654435761 + l * 97) % 4 for l, n in enumerate(sizes) for i in range(first, n, stride)], np.uint8)


def test_sort_free_placement_equals_key_merge():
    """pm_exchange_photons' layout arithmetic (csrc/pm_comm.cu), stated in numpy: count bytes + two prefix sums give the
    same array as the keyed merge, for ragged lights (fewer paths than ranks, empty lights) and any world size."""
    mg = graft.load_package_module("multigpu")
    for sizes in ([40, 13, 27], [0, 5, 0], [1], [3, 2, 1, 0, 9], []):
        full_rec, _ = fake_emit(sizes, 0, 1)
        for world in (1, 2, 3, 5, 8):
            shards = [fake_emit(sizes, r, world)[0] for r in range(world)]
            counts = [fake_counts(sizes, r, world) for r in range(world)]
            assert [len(c) for c in counts] == [mg.rank_path_count(sizes, r, world) for r in range(world)]
            got = mg.place_by_counts(counts, shards, sizes)
            assert np.array_equal(got, full_rec.reshape(-1, 40))
    with pytest.raises(AssertionError):
        mg.place_by_counts([np.zeros(3, np.uint8)], [np.zeros((0, 40), np.uint8)], [4])


def test_shard_arithmetic():
    mg = graft.load_package_module("multigpu")
    assert [mg.shard_range(10, r, 4) for r in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert mg.shard_range(0, 0, 2) == (0, 0)
    for total in (1, 7, 64, 1001):
        for world in (1, 2, 3, 8):
            spans = [mg.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    sizes = mg.light_launch_sizes(_lights(), 200000)
    assert sizes == [100000, 25000, 75000]  # int(casted / 20) = 10000 per watt (hostCode.cu:102-110)
    for world in (1, 2, 3, 8):
        assert sum(mg.rank_path_count(sizes, r, world) for r in range(world)) == sum(sizes)
    assert mg.light_launch_sizes(_lights(), 7) == [0, 0, 0]  # truncation to 0 photons per watt


def test_merge_restores_single_gpu_order():
    mg = graft.load_package_module("multigpu")
    sizes = [40, 13, 27]
    full_rec, full_keys = fake_emit(sizes, 0, 1)
    for world in (2, 3, 5):
        shards = [fake_emit(sizes, r, world) for r in range(world)]
        rec, keys = mg.merge_shards([s[0] for s in shards], [s[1] for s in shards])
        assert np.array_equal(keys, full_keys) and np.array_equal(rec, full_rec)


def _worker(rank, world, port, sizes, out_dir):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mg = graft.load_package_module("multigpu")
    rec, keys = fake_emit(sizes, rank, world)
    all_rec, all_keys = mg.allgather_photons(torch.from_numpy(rec.copy()).reshape(-1), torch.from_numpy(keys.view(np.int64).copy()))
    np.save(os.path.join(out_dir, f"rec{rank}.npy"), all_rec.numpy())
    np.save(os.path.join(out_dir, f"key{rank}.npy"), all_keys.numpy())
    # the tensor-resident exchange (what tools/c4_run.py runs over NCCL) gives the same array
    x_rec, x_keys, counts = mg.exchange_photons(torch.from_numpy(rec.copy()), torch.from_numpy(keys.view(np.int64).copy()))
    assert torch.equal(x_rec, all_rec.reshape(-1, 40)) and torch.equal(x_keys, all_keys) and counts[rank] == len(keys)
    # the library's sort-free exchange, same collectives over gloo: padded count rows and record rows, then placement
    cnt = fake_counts(sizes, rank, world)
    pmax = max(mg.rank_path_count(sizes, r, world) for r in range(world))
    row = torch.zeros(pmax, dtype=torch.uint8)
    row[: len(cnt)] = torch.from_numpy(cnt.copy())
    rows = [torch.zeros_like(row) for _ in range(world)]
    dist.all_gather(rows, row)
    stored = [int(r.sum()) for r in rows]
    pad = torch.zeros((max(stored), 40), dtype=torch.uint8)
    pad[: len(rec)] = torch.from_numpy(rec.copy())
    pads = [torch.zeros_like(pad) for _ in range(world)]
    dist.all_gather(pads, pad)
    placed = mg.place_by_counts([rows[r][: mg.rank_path_count(sizes, r, world)].numpy() for r in range(world)],
                                [pads[r][: stored[r]].numpy() for r in range(world)], sizes)
    assert np.array_equal(placed, all_rec.numpy().reshape(-1, 40))
    # progressive mode: one sum of (flux.rgb, count) per pass over disjoint photon shards
    ppm = graft.load_package_module("ppm")
    rng = np.random.default_rng(100 + rank)
    flux = rng.random((50, 3), dtype=np.float32)
    count = rng.integers(0, 2000, 50).astype(np.int32)
    f_sum, c_sum = ppm._allreduce(flux, count, None, None)
    np.save(os.path.join(out_dir, f"ppm{rank}.npy"), np.concatenate([f_sum, c_sum[:, None].astype(np.float32)], axis=1))
    # stage 3: disjoint pixel ranges merged by a sum
    W, H = 37, 11
    b, e = mg.shard_range(W * H, rank, world)
    fb = np.zeros(W * H, np.int64)
    fb[b:e] = (np.arange(b, e) * 2654435761) % (2**32)
    t = torch.from_numpy(fb)
    dist.all_reduce(t)
    np.save(os.path.join(out_dir, f"fb{rank}.npy"), t.numpy())
    dist.destroy_process_group()


def test_gloo_world2_allgather_and_tiles(tmp_path):
    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    sizes = [50, 0, 21]
    world = 2
    mp.spawn(_worker, args=(world, port, sizes, str(tmp_path)), nprocs=world, join=True)
    full_rec, full_keys = fake_emit(sizes, 0, 1)
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / f"key{r}.npy").view(np.uint64), full_keys)
        assert np.array_equal(np.load(tmp_path / f"rec{r}.npy"), full_rec)
        want = (np.arange(37 * 11) * 2654435761) % (2**32)
        assert np.array_equal(np.load(tmp_path / f"fb{r}.npy"), want)
    parts = [np.random.default_rng(100 + r) for r in range(world)]
    flux = sum(g.random((50, 3), dtype=np.float32).astype(np.float64) for g in parts)
    parts = [np.random.default_rng(100 + r) for r in range(world)]
    count = 0
    for g in parts:
        g.random((50, 3), dtype=np.float32)
        count = count + g.integers(0, 2000, 50)
    for r in range(world):
        got = np.load(tmp_path / f"ppm{r}.npy")
        assert np.array_equal(got[:, 3], count.astype(np.float32)) and np.allclose(got[:, :3], flux, rtol=1e-6)

"""GPU test of the sharded stage 1 on ONE GPU: the shards of G emulated ranks (first_photon = r,
photon_stride = G), merged by key, equal the single-rank emission record for record; pixel ranges
merge into the full framebuffer (covered in test_gpu_render.py)."""
import numpy as np
import pytest

import __graft_entry__ as graft
from helpers import load_scene

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("caustics", [False, True])
@pytest.mark.parametrize("world", [2, 3])
def test_sharded_emission_merges_to_single_gpu_order(pkg, ctx, world, caustics):
    mg = graft.load_package_module("multigpu")
    sc = load_scene("sphere" if caustics else "cornell_box")
    scene = pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])
    casted, depth = 30000, 6
    full, n, paths, _ = scene.emit_photons(sc["lights"], casted, depth, caustics)
    full_k, keys_k, paths_k, _ = scene.emit_photons_keyed(sc["lights"], casted, depth, caustics)
    assert full_k.tobytes() == full.tobytes() and (np.diff(keys_k.astype(np.int64)) > 0).all()  # keys strictly increase
    sizes = mg.light_launch_sizes(sc["lights"], casted)
    assert paths == sum(sizes) == paths_k
    recs, keys = [], []
    for r in range(world):
        rec, key, p, _ = scene.emit_photons_keyed(sc["lights"], casted, depth, caustics, first_photon=r, photon_stride=world)
        assert p == mg.rank_path_count(sizes, r, world)
        recs.append(rec.view(np.uint8).reshape(-1, 40))
        keys.append(key)
    merged, mk = mg.merge_shards(recs, keys)
    assert np.array_equal(mk, keys_k)
    assert merged.tobytes() == full.tobytes()


@pytest.fixture(params=[False, True], ids=["lists", "carried-keys"])
def carried(request, pkg):
    """both variants of the level pass: the default below 2^24 photons, and the large-map one forced on"""
    if request.param:
        pkg.lib().pm_test_build_carry_from(1)
    yield request.param
    pkg.lib().pm_test_build_carry_from(0)


@pytest.mark.parametrize("n,world", [(300000, 2), (300000, 8), (1200001, 4), (5000, 4), (3, 2)])
def test_sharded_build_equals_single_gpu_build(pkg, ctx, n, world, carried):
    """pm_map_build_shard for every emulated rank of `world` on ONE GPU, node tables summed as the all-reduce would,
    pm_map_build_finish: the tree array (split_dim included) and the bounds equal pm_map_build's byte for byte —
    unique coordinates and a tie-heavy lattice; trees too small to shard take the whole-build path."""
    import torch
    from helpers import dup_cloud, unique_cloud

    for cloud in (unique_cloud(n, 31), dup_cloud(n, 32, 5)):
        want = torch.from_numpy(cloud.copy().view(np.uint8)).cuda()
        ref = pkg.PhotonMap.build(ctx, want)
        elems, dims = [], []
        src = torch.from_numpy(cloud.copy().view(np.uint8)).cuda()
        before = src.clone()
        bounds = np.zeros(6, np.float32)
        for r in range(world):
            elem = torch.empty(n, dtype=torch.int32, device="cuda")
            dim = torch.zeros((n + 3) // 4 * 4, dtype=torch.uint8, device="cuda")
            pkg._check(pkg.lib().pm_map_build_shard(ctx.handle, src.data_ptr(), n, r, world, elem.data_ptr(), dim.data_ptr(),
                                                    bounds.ctypes.data))
            elems.append(elem)
            dims.append(dim.view(torch.int32).clone())
        assert torch.equal(src, before)  # the shard pass does not touch the photons
        elem, dim = sum(elems[1:], elems[0]), sum(dims[1:], dims[0])
        import ctypes

        h = ctypes.c_void_p()
        pkg._check(pkg.lib().pm_map_build_finish(ctx.handle, src.data_ptr(), n, elem.data_ptr(), dim.data_ptr(), bounds.ctypes.data,
                                                 ctypes.byref(h)))
        got = pkg.PhotonMap(ctx, h, n, bounds, src)
        assert torch.equal(src, want) and np.array_equal(bounds, ref.bounds)
        q = torch.rand((500, 3), device="cuda") * float(cloud["pos"].max())
        a, b = got.knn(q, 8, 100.0), ref.knn(q, 8, 100.0)
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
        got.close()
        ref.close()


@pytest.mark.parametrize("caustics,power", [(False, 1.0), (True, 0.5)])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_sortfree_exchange_places_shards_in_single_gpu_order(pkg, ctx, world, caustics, power):
    """pm_emit_photons_counted per emulated rank + pm_exchange_place (the single-GPU part of pm_exchange_photons: two
    prefix sums over the gathered per-path count bytes + the placement kernel) equal the single-rank emission
    converted by pm_photons_to_maps, byte for byte — no keys, no sort."""
    import ctypes

    import torch

    L = pkg.lib()
    sc = load_scene("sphere" if caustics else "cornell_box")
    scene = pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])
    lights = np.ascontiguousarray(sc["lights"], pkg.LIGHT)
    casted, depth = 50001, 7
    full, n, paths, _ = scene.emit_photons(lights, casted, depth, caustics, seed=3, rng_mode=pkg.RNG_COUNTER)
    want = np.zeros(n, pkg.PHOTON44)
    pkg._check(L.pm_photons_to_maps(ctx.handle, full.ctypes.data, n, None, 0, want.ctypes.data, None))
    want["power"] = power
    mg = graft.load_package_module("multigpu")
    sizes = mg.light_launch_sizes(lights, casted)
    pmax = max(mg.rank_path_count(sizes, r, world) for r in range(world))
    pitch = (pmax + 15) // 16 * 16 + 16
    cap_r = casted * depth // world + 64
    counts_all = torch.zeros((world, pitch), dtype=torch.uint8, device="cuda")
    shards = torch.zeros((world, cap_r, 40), dtype=torch.uint8, device="cuda")
    total_local = 0
    for r in range(world):
        cnt, npaths, nseg = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int64(0)
        pkg._check(L.pm_emit_photons_counted(ctx.handle, scene.handle, lights.ctypes.data, None, len(lights), casted, depth,
                                             1 if caustics else 0, 3, pkg.RNG_COUNTER, r, world, shards[r].data_ptr(),
                                             counts_all[r].data_ptr(), cap_r, ctypes.byref(cnt), ctypes.byref(npaths),
                                             ctypes.byref(nseg)))
        assert npaths.value == mg.rank_path_count(sizes, r, world)
        assert int(counts_all[r].sum().item()) == cnt.value
        total_local += cnt.value
    assert total_local == n
    out = torch.zeros((n + 5, 44), dtype=torch.uint8, device="cuda")
    total = ctypes.c_int64(0)
    pkg._check(L.pm_exchange_place(ctx.handle, world, lights.ctypes.data, len(lights), casted, counts_all.data_ptr(), pitch,
                                   shards.data_ptr(), cap_r, power, out.data_ptr(), n + 5, ctypes.byref(total)))
    assert total.value == n
    got = out[:n].cpu().numpy().reshape(-1).view(pkg.PHOTON44)
    assert got.tobytes() == want.tobytes()
    assert int(out[n:].sum().item()) == 0  # nothing written past the end
    # too small an output is reported, not overrun
    rc = L.pm_exchange_place(ctx.handle, world, lights.ctypes.data, len(lights), casted, counts_all.data_ptr(), pitch,
                             shards.data_ptr(), cap_r, power, out.data_ptr(), n - 1, ctypes.byref(total))
    assert rc == -5 and total.value == n  # PM_ERR_CAPACITY


def test_ppm_update_kernel_equals_numpy_update(pkg, ctx):
    """pm_ppm_update (one rank: no collective) against ppm.ppm_update, bit for bit."""
    import ctypes

    import torch

    ppm = graft.load_package_module("ppm")
    rng = np.random.default_rng(12)
    npx = 50000
    N = (rng.random(npx, dtype=np.float32) * 40).astype(np.float32)
    N[::7] = 0
    r2 = (rng.random(npx, dtype=np.float32) + np.float32(0.01)).astype(np.float32)
    tau = rng.random((npx, 3), dtype=np.float32)
    flux = rng.random((npx, 3), dtype=np.float32)
    count = rng.integers(0, 30, npx).astype(np.int32)
    count[::5] = 0
    flux[count == 0] = 0
    dN, dr2, dtau = (torch.from_numpy(a.copy()).cuda() for a in (N, r2, tau))
    dflux, dcount = torch.from_numpy(flux).cuda(), torch.from_numpy(count).cuda()
    pkg._check(pkg.lib().pm_ppm_update(ctx.handle, None, npx, dflux.data_ptr(), dcount.data_ptr(), 0.7, dN.data_ptr(),
                                       dr2.data_ptr(), dtau.data_ptr(), None))
    ppm.ppm_update(N, r2, tau, flux, count, 0.7)
    assert np.array_equal(dN.cpu().numpy(), N) and np.array_equal(dr2.cpu().numpy(), r2)
    assert np.array_equal(dtau.cpu().numpy(), tau)

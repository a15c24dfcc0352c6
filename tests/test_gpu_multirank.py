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


@pytest.mark.parametrize("n,world", [(300000, 2), (300000, 8), (1200001, 4), (5000, 4), (3, 2)])
def test_sharded_build_equals_single_gpu_build(pkg, ctx, n, world):
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

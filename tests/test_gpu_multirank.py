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

"""GPU parity tests for the scene BVH (hit ids bit-exact vs the brute-force oracle) and stage 1
(photon emission + walk, byte-exact vs the oracle restatement of photon-mapping/cuda/deviceCode.cu)."""
import numpy as np
import pytest

from helpers import (PHOTON40, PHOTON44, load_scene, oracle_emit, oracle_occluded, oracle_trace, random_rays,
                     soup_scene, vp)

pytestmark = pytest.mark.gpu


def make_scene(pkg, ctx, sc):
    return pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])


def scene_box(sc):
    v = sc["verts"][sc["tris"].reshape(-1)]
    return v.min(0), v.max(0)


@pytest.mark.parametrize("name,nrays", [("cornell_box", 200000), ("sphere", 20000)])
def test_trace_closest_matches_bruteforce_on_reference_assets(pkg, ctx, orc, name, nrays):
    sc = load_scene(name)
    s = make_scene(pkg, ctx, sc)
    lo, hi = scene_box(sc)
    rays = random_rays(nrays, 11, lo - 1, hi + 1)
    # a block of rays leaving the light positions (what stage 1 actually shoots)
    rays["org"][: nrays // 4] = sc["lights"]["pos"][0]
    got = s.trace_closest(rays)
    want = oracle_trace(orc, sc, rays)
    assert (want["prim"] >= 0).mean() > 0.5
    assert np.array_equal(got["prim"], want["prim"])            # hit ids bit-exact
    assert got["t"].tobytes() == want["t"].tobytes()            # and so is t
    occ = s.trace_occluded(rays)
    assert np.array_equal(occ, oracle_occluded(orc, sc, rays))


def test_trace_closest_dragon_box(pkg, ctx, orc):
    """assets/models/dragon/dragon-box.glb (91 226 triangles, the reference's largest asset) through the LBVH."""
    sc = load_scene("dragon_box")
    assert len(sc["tris"]) == 91226
    s = make_scene(pkg, ctx, sc)
    lo, hi = scene_box(sc)
    rays = random_rays(12000, 23, lo - 0.5, hi + 0.5)
    rays["org"][:3000] = sc["lights"]["pos"][0]  # what stage 1 shoots
    # a block aimed at the dragon itself (the walls take most random rays)
    dv = sc["verts"][sc["tris"][sc["tri_mesh"] == 0].reshape(-1)]
    tgt = dv[np.random.default_rng(5).integers(0, len(dv), 3000)]
    d = tgt - rays["org"][3000:6000]
    rays["dir"][3000:6000] = d / np.linalg.norm(d, axis=1, keepdims=True)
    got = s.trace_closest(rays)
    want = oracle_trace(orc, sc, rays)
    assert (sc["tri_mesh"][want["prim"][want["prim"] >= 0]] == 0).sum() > 2000  # the dragon is hit
    assert np.array_equal(got["prim"], want["prim"])
    assert got["t"].tobytes() == want["t"].tobytes()
    assert np.array_equal(s.trace_occluded(rays), oracle_occluded(orc, sc, rays))


def test_trace_closest_c4_room_1m_triangles(pkg, ctx, orc):
    """The procedural 1 000 012-triangle room of C4 (photon-mapping_b200/scenes.py) against brute force."""
    import __graft_entry__ as graft

    sc = graft.load_package_module("scenes").c4_room()
    assert len(sc["tris"]) == 1_000_012
    s = make_scene(pkg, ctx, sc)
    rays = random_rays(2400, 31, [-20, 0, -20], [20, 40, 20])
    rays["org"][:600] = sc["lights"]["pos"][0]
    rays["org"][600:1200] = sc["lights"]["pos"][1]
    got = s.trace_closest(rays)
    want = oracle_trace(orc, sc, rays)
    assert (want["prim"] >= 0).mean() > 0.9 and (want["prim"] < 1_000_000).sum() > 300  # walls and the sphere
    assert np.array_equal(got["prim"], want["prim"])
    assert got["t"].tobytes() == want["t"].tobytes()
    assert np.array_equal(s.trace_occluded(rays), oracle_occluded(orc, sc, rays))


@pytest.mark.parametrize("nt", [1, 2, 3, 17, 1000, 4096, 4097, 20000])  # <= 4096: host SAH hierarchy, above: LBVH
def test_trace_closest_soup(pkg, ctx, orc, nt):
    sc = soup_scene(nt, nt)
    s = make_scene(pkg, ctx, sc)
    rays = random_rays(20000 if nt <= 1000 else 4000, nt + 1, [-1, -1, -1], [11, 11, 11])
    rays["tmax"][::3] = 6.0   # finite tmax
    got = s.trace_closest(rays)
    want = oracle_trace(orc, sc, rays)
    assert np.array_equal(got["prim"], want["prim"])
    assert got["t"].tobytes() == want["t"].tobytes()
    assert np.array_equal(s.trace_occluded(rays), oracle_occluded(orc, sc, rays))


def test_trace_nested_triangles_deep_sah_chain(pkg, ctx, orc):
    """Concentric coplanar-ish triangles of growing size: the SAH sweep peels one triangle per level, so the
    builder's depth guard (median splits below level 44) has to keep the ray stack within its 64 entries."""
    n = 300
    k = np.arange(1, n + 1, dtype=np.float32)
    ang = np.array([0.0, 2.1, 4.2], np.float32)
    verts = np.stack([np.stack([k * np.cos(a), k * np.sin(a), 0.01 * k], 1) for a in ang], 1).reshape(-1, 3).astype(np.float32)
    sc = {"verts": verts, "tris": np.arange(3 * n, dtype=np.int32).reshape(n, 3), "tri_mesh": np.zeros(n, np.int32),
          "mats": np.zeros(1, pkg.MATERIAL)}
    s = make_scene(pkg, ctx, sc)
    rays = random_rays(20000, 77, [-50, -50, -20], [50, 50, 20])
    got, want = s.trace_closest(rays), oracle_trace(orc, sc, rays)
    assert (want["prim"] >= 0).sum() > 1000
    assert np.array_equal(got["prim"], want["prim"]) and got["t"].tobytes() == want["t"].tobytes()
    assert np.array_equal(s.trace_occluded(rays), oracle_occluded(orc, sc, rays))


def test_trace_duplicate_triangles_tie_to_lowest_id(pkg, ctx, orc):
    sc = soup_scene(50, 5)
    # every triangle twice: identical t, the lower id must win
    sc["verts"] = np.concatenate([sc["verts"], sc["verts"]])
    sc["tris"] = np.concatenate([sc["tris"] + 150, sc["tris"]]).astype(np.int32)
    sc["tri_mesh"] = np.concatenate([sc["tri_mesh"], sc["tri_mesh"]])
    s = make_scene(pkg, ctx, sc)
    rays = random_rays(20000, 6, [-1, -1, -1], [11, 11, 11])
    got, want = s.trace_closest(rays), oracle_trace(orc, sc, rays)
    assert (want["prim"] >= 0).sum() > 100 and (want["prim"][want["prim"] >= 0] < 50).all()
    assert np.array_equal(got["prim"], want["prim"]) and got["t"].tobytes() == want["t"].tobytes()


def test_empty_scene(pkg, ctx):
    sc = soup_scene(1, 1)
    s = pkg.Scene(ctx, np.zeros((0, 3), np.float32), np.zeros((0, 3), np.int32), np.zeros(0, np.int32), sc["mats"])
    rays = random_rays(100, 1, [0, 0, 0], [1, 1, 1])
    h = s.trace_closest(rays)
    assert (h["prim"] == -1).all() and np.array_equal(h["t"], rays["tmax"])
    ph, cnt, npaths, nseg = s.emit_photons(sc["lights"], 1000, 5, False)
    assert cnt == 0 and npaths == 990 and nseg == 990   # int(1000/15)=66 per watt -> 660 + 330 paths, all miss


@pytest.mark.parametrize("caustics", [False, True])
@pytest.mark.parametrize("scene_name", ["cornell_box", "sphere", "soup", "soup_6000"])
def test_emit_matches_oracle_lcg(pkg, ctx, orc, scene_name, caustics):
    """soup_6000 is above the 4 096 triangles from which the walk uses the vote-driven warp traversal (k_walk<true>) on
    a Morton-built hierarchy; the others take the while-while loop on the host-built SAH hierarchy"""
    if scene_name == "soup":
        sc = soup_scene(3000, 9, extent=6.0, size=2.5)
    elif scene_name == "soup_6000":
        sc = soup_scene(6000, 10, extent=6.0, size=1.8)
    else:
        sc = load_scene(scene_name)
    casted, depth = (20000, 10) if scene_name != "sphere" else (3000, 6)
    s = make_scene(pkg, ctx, sc)
    got, cnt, npaths, nseg = s.emit_photons(sc["lights"], casted, depth, caustics)
    want, o_paths, o_seg = oracle_emit(orc, sc, casted, depth, caustics)
    assert (npaths, nseg) == (o_paths, o_seg)
    assert cnt == len(want)
    if scene_name == "cornell_box" and not caustics:
        assert cnt > casted // 4          # the walk really stores photons
    assert got.tobytes() == want.tobytes()    # same photons, same order, every byte


def test_emit_counter_rng_and_sharding(pkg, ctx, orc):
    sc = load_scene("cornell_box")
    s = make_scene(pkg, ctx, sc)
    full, cnt, npaths, nseg = s.emit_photons(sc["lights"], 10000, 8, False, seed=77, rng_mode=pkg.RNG_COUNTER)
    want, o_paths, o_seg = oracle_emit(orc, sc, 10000, 8, False, seed=77, rng_mode=1)
    assert full.tobytes() == want.tobytes() and (npaths, nseg) == (o_paths, o_seg)
    other, _, _, _ = s.emit_photons(sc["lights"], 10000, 8, False, seed=78, rng_mode=pkg.RNG_COUNTER)
    assert other.tobytes() != full.tobytes()
    # shard by launch index over 3 ranks: the union is the full set, each shard matches the oracle
    parts, paths = [], 0
    for r in range(3):
        p, c, npth, _ = s.emit_photons(sc["lights"], 10000, 8, False, seed=77, rng_mode=pkg.RNG_COUNTER,
                                      first_photon=r, photon_stride=3)
        w, _, _ = oracle_emit(orc, sc, 10000, 8, False, seed=77, rng_mode=1, first=r, stride=3)
        assert p.tobytes() == w.tobytes()
        parts.append(p.copy())
        paths += npth
    assert paths == npaths
    allp = np.concatenate(parts)
    assert sorted(allp.tobytes()[i:i + 40] for i in range(0, allp.nbytes, 40)) == \
        sorted(full.tobytes()[i:i + 40] for i in range(0, full.nbytes, 40))


def test_emit_capacity_and_device_output(pkg, ctx, orc):
    import torch

    sc = load_scene("cornell_box")
    s = make_scene(pkg, ctx, sc)
    want, _, _ = oracle_emit(orc, sc, 5000, 10, False)
    with pytest.raises(pkg.PmError) as e:
        s.emit_photons(sc["lights"], 5000, 10, False, capacity=100)
    assert e.value.code == -5
    d_out = torch.zeros(5000 * 10 * 40, dtype=torch.uint8, device="cuda")
    _, cnt, _, _ = s.emit_photons(sc["lights"], 5000, 10, False, out=d_out, capacity=5000 * 10)
    assert cnt == len(want)
    assert d_out[: cnt * 40].cpu().numpy().tobytes() == want.tobytes()
    # depth 1 stores nothing (i > 0 rule), depth 0 traces nothing
    _, c1, _, s1 = s.emit_photons(sc["lights"], 5000, 1, False)
    assert c1 == 0 and s1 == 5000
    _, c0, _, s0 = s.emit_photons(sc["lights"], 5000, 0, False)
    assert c0 == 0 and s0 == 0


def test_photons_to_maps(pkg, ctx, orc):
    sc = load_scene("cornell_box")
    s = make_scene(pkg, ctx, sc)
    g, _, _, _ = s.emit_photons(sc["lights"], 4000, 10, False)
    c, _, _, _ = s.emit_photons(sc["lights"], 4000, 10, True)
    og, oc = pkg.photons_to_maps(ctx, g, c)
    wg, wc = np.zeros(len(g) + len(c), PHOTON44), np.zeros(len(c), PHOTON44)
    orc.orc_photons_to_maps(vp(np.ascontiguousarray(g)), len(g), vp(np.ascontiguousarray(c)), len(c), vp(wg), vp(wc))
    assert og.tobytes() == wg.tobytes() and oc.tobytes() == wc.tobytes()
    assert (og["power"][: len(g)] == 1.0).all() and (og["power"][len(g):] == 0.5).all()


@pytest.mark.parametrize("rng_mode", [0, 1])
def test_square_light_matches_oracle(pkg, ctx, orc, rng_mode):
    """Square emitters (the reference declares LightSource::normal / side_length,
    common/src/world.h:17-20, and never reads them): bit-exact against the oracle's restatement of OUR definition."""
    from helpers import LIGHT_SHAPE
    sc = load_scene("cornell_box")
    s = make_scene(pkg, ctx, sc)
    shapes = np.zeros(len(sc["lights"]), LIGHT_SHAPE)
    shapes[0] = (1, (0.0, -1.0, 0.0), 6.0)        # a ceiling panel shining down; the second light stays a point
    got, cnt, npaths, nseg = s.emit_photons(sc["lights"], 20000, 8, False, seed=5, rng_mode=rng_mode, shapes=shapes)
    want, o_paths, o_seg = oracle_emit(orc, sc, 20000, 8, False, seed=5, rng_mode=rng_mode, shapes=shapes)
    assert (npaths, nseg) == (o_paths, o_seg) and cnt == len(want) and cnt > 2000
    assert got.tobytes() == want.tobytes()
    # all-point shapes reproduce the unshaped call; a square light changes the result
    pts, c2, _, _ = s.emit_photons(sc["lights"], 20000, 8, False, seed=5, rng_mode=rng_mode, shapes=np.zeros(len(sc["lights"]), LIGHT_SHAPE))
    plain, c3, _, _ = s.emit_photons(sc["lights"], 20000, 8, False, seed=5, rng_mode=rng_mode)
    assert pts[:c2].tobytes() == plain[:c3].tobytes() and plain[:c3].tobytes() != got.tobytes()
    bad = shapes.copy(); bad[0]["side_length"] = 0
    with pytest.raises(pkg.PmError):
        s.emit_photons(sc["lights"], 100, 4, False, shapes=bad)

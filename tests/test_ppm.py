"""Progressive (fixed-radius) photon mapping — BASELINE config 5. No reference implementation exists, so parity is
self-consistency: the host loop's update arithmetic (CPU), and the GPU passes against the same loop driven by the
oracle (oracle photon walk with the counter RNG + brute-force fixed-radius gather)."""
import ctypes
from ctypes import c_int64, c_void_p

import numpy as np
import pytest

import __graft_entry__ as graft
from helpers import PHOTON44, load_scene, oracle_emit, vp


def test_ppm_update_arithmetic():
    ppm = graft.load_package_module("ppm")
    N = np.array([0, 10, 10, 5], np.float32)
    r2 = np.array([4, 4, 4, 1], np.float32)
    tau = np.ones((4, 3), np.float32)
    flux = np.full((4, 3), 2, np.float32)
    flux[2] = 0
    count = np.array([8, 10, 0, 5], np.int32)
    ppm.ppm_update(N, r2, tau, flux, count, 0.7)
    assert np.allclose(N, [5.6, 17, 10, 8.5])
    assert np.allclose(r2, [4 * 5.6 / 8, 4 * 17 / 20, 4, 8.5 / 10])
    assert np.allclose(tau[:, 0], [3 * 5.6 / 8, 3 * 17 / 20, 1, 3 * 8.5 / 10])
    assert (r2 <= [4, 4, 4, 1]).all()  # radii never grow


def test_paths_launched_follows_the_per_watt_truncation():
    """photon-mapping/src/hostCode.cu:86,102-110: int(power * int(casted / sum(power))) per light."""
    ppm = graft.load_package_module("ppm")
    LIGHT = np.dtype([("pos", "<f4", 3), ("rgb", "<f4", 3), ("power", "<f8")])
    lights = np.zeros(3, LIGHT)
    lights["power"] = [10.0, 2.5, 0.3]
    assert ppm.paths_launched(lights, 1000) == int(10.0 * 78) + int(2.5 * 78) + int(0.3 * 78) == 998
    assert ppm.paths_launched(lights[:1], 999) == 990 and ppm.paths_launched(lights[:0], 5) == 0


@pytest.mark.gpu
def test_progressive_passes_match_oracle_loop(pkg, ctx, orc):
    ppm = graft.load_package_module("ppm")
    sc = load_scene("cornell_box")
    scene = pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])
    W, H, passes, per_pass, r0, alpha, depth = 48, 32, 3, 20000, 3.0, 0.7, 6
    cam = pkg.make_camera((80.0, 30.0, 0.0), (10.0, 20.0, 0.0), (0.0, 1.0, 0.0), 0.87, W, H)
    got = ppm.progressive_render(pkg, ctx, scene, sc["tri_mesh"], sc["mats"], sc["lights"], cam, W, H, passes, per_pass, r0, alpha,
                                 depth, seed=5)
    # the same loop on the oracle
    rays = ppm.camera_rays(pkg, cam, W, H)
    pos, brdf, albedo = ppm.hit_points(pkg, scene, sc["tri_mesh"], sc["mats"], rays)
    assert (brdf > 0).mean() > 0.5
    N = np.zeros(W * H, np.float32)
    r2 = np.full(W * H, np.float32(r0) ** 2, np.float32)
    tau = np.zeros((W * H, 3), np.float32)
    orc.orc_gather_radius.argtypes = [c_void_p, ctypes.c_int, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]
    for p in range(passes):
        g40, _, _ = oracle_emit(orc, sc, per_pass, depth, False, seed=5 + p, rng_mode=1)
        g44 = np.zeros(len(g40), PHOTON44)
        g44["pos"], g44["dir"], g44["color"], g44["power"] = g40["pos"], g40["dir"], g40["color"], 1.0
        flux = np.zeros((W * H, 3), np.float32)
        cnt = np.zeros(W * H, np.int32)
        orc.orc_gather_radius(vp(g44), len(g44), vp(pos), vp(brdf), vp(r2), W * H, vp(flux), vp(cnt))
        ppm.ppm_update(N, r2, tau, flux, cnt, alpha)
    assert np.array_equal(got["N"], N) and np.array_equal(got["r2"], r2)  # counts are exact, so are N and r²
    assert np.allclose(got["tau"], tau, rtol=2e-4, atol=1e-6)
    assert got["emitted"] == passes * ppm.paths_launched(sc["lights"], per_pass) and np.isfinite(got["radiance"]).all() and got["radiance"].max() > 0
    assert (r2 < np.float32(r0) ** 2).any()


@pytest.mark.gpu
def test_device_resident_loop_equals_host_loop(pkg, ctx):
    """progressive_render_device keeps photons, maps and per-pixel state in HBM; on one GPU its N, r² and tau equal
    the host loop's bit for bit (same C-ABI calls, the update restated in torch float32)."""
    ppm = graft.load_package_module("ppm")
    sc = load_scene("cornell_box")
    scene = pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])
    W, H, passes, per_pass, r0, alpha, depth = 64, 40, 4, 30000, 2.5, 0.7, 6
    cam = pkg.make_camera((80.0, 30.0, 0.0), (10.0, 20.0, 0.0), (0.0, 1.0, 0.0), 0.87, W, H)
    host = ppm.progressive_render(pkg, ctx, scene, sc["tri_mesh"], sc["mats"], sc["lights"], cam, W, H, passes, per_pass, r0, alpha,
                                  depth, seed=9)
    stats = []
    dev = ppm.progressive_render_device(pkg, ctx, scene, sc["tri_mesh"], sc["mats"], sc["lights"], cam, W, H, passes, per_pass, r0,
                                        alpha, depth, seed=9, stats=stats)
    assert np.array_equal(dev["N"], host["N"]) and np.array_equal(dev["r2"], host["r2"])
    assert np.array_equal(dev["tau"], host["tau"]) and np.array_equal(dev["radiance"], host["radiance"])
    assert len(stats) == passes and all(s["stored"] > 0 for s in stats)
    assert dev["emitted"] == sum(s["paths"] for s in stats) == passes * ppm.paths_launched(sc["lights"], per_pass)


@pytest.mark.gpu
def test_cells_only_map_counts_and_rejects_knn(pkg, ctx, orc):
    """pm_map_build_cells: fixed-radius gathers from a map without a kd-tree — counts exact and flux within 1e-4 of the
    brute-force oracle for radii from a fraction of a cell to the whole cloud (blocks grow to the whole grid); kNN
    entry points refuse the map; the progressive loop on cells-only maps reproduces N and r² of the kd-tree loop."""
    from helpers import unique_cloud

    n, nq = 30000, 1500
    cloud = unique_cloud(n, 21)
    ext = n * 2.0**-8
    rng = np.random.default_rng(6)
    q = (rng.random((nq, 3), dtype=np.float32) * np.float32(1.2 * ext) - np.float32(0.1 * ext)).astype(np.float32)
    brdf = (rng.random(nq, dtype=np.float32) * 0.9 + 0.1).astype(np.float32)
    brdf[::13] = 0.0
    gap = ext / n ** (1 / 3)
    r2 = ((rng.random(nq, dtype=np.float32) * 6.0 + 0.2) * np.float32(gap)) ** 2
    r2[::97] = np.float32(4 * ext * ext)  # the whole cloud
    want_f = np.zeros((nq, 3), np.float32)
    want_c = np.zeros(nq, np.int32)
    orc.orc_gather_radius.argtypes = [c_void_p, ctypes.c_int, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]
    orc.orc_gather_radius(vp(cloud), n, vp(q), vp(brdf), vp(r2), nq, vp(want_f), vp(want_c))
    for cell in (0.7 * gap, 3.0 * gap):
        m = pkg.PhotonMap.build_cells(ctx, cloud, float(cell))
        flux, cnt = m.gather_radius(q, brdf, r2)
        assert np.array_equal(cnt, want_c) and want_c.max() == n
        assert np.allclose(flux, want_f, rtol=1e-4, atol=1e-6)
        with pytest.raises(pkg.PmError):
            m.knn(q, 8)
        with pytest.raises(pkg.PmError):
            m.gather(q, brdf, 8)
        m.close()
    with pytest.raises(pkg.PmError):
        pkg.PhotonMap.build_cells(ctx, cloud, 0.0)

    ppm = graft.load_package_module("ppm")
    sc = load_scene("cornell_box")
    scene = pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])
    W, H = 64, 40
    cam = pkg.make_camera((80.0, 30.0, 0.0), (10.0, 20.0, 0.0), (0.0, 1.0, 0.0), 0.87, W, H)
    a = ppm.progressive_render_device(pkg, ctx, scene, sc["tri_mesh"], sc["mats"], sc["lights"], cam, W, H, 4, 30000, 2.5, 0.7, 6, seed=9)
    b = ppm.progressive_render_device(pkg, ctx, scene, sc["tri_mesh"], sc["mats"], sc["lights"], cam, W, H, 4, 30000, 2.5, 0.7, 6, seed=9,
                                      cells=True)
    assert np.array_equal(a["N"], b["N"]) and np.array_equal(a["r2"], b["r2"])
    assert np.allclose(a["tau"], b["tau"], rtol=2e-4, atol=1e-6)

"""GPU parity tests for stage 3 (pm_render) against oracle/orc_render.cpp on the reference's Cornell box."""
import numpy as np
import pytest

from helpers import PHOTON44, load_scene, oracle_render, vp

pytestmark = pytest.mark.gpu


def camera(pkg, W, H):
    return pkg.make_camera([80.0, 30.0, 0.0], [10.0, 20.0, 0.0], [0.0, 1.0, 0.0], 0.87, W, H)


@pytest.fixture(scope="module")
def cornell(pkg, ctx):
    sc = load_scene("cornell_box")
    s = pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])
    g40, _, _, _ = s.emit_photons(sc["lights"], 20000, 10, False)
    c40, _, _, _ = s.emit_photons(sc["lights"], 10000, 10, True)
    g44, c44 = pkg.photons_to_maps(ctx, g40, c40)
    gmap = pkg.PhotonMap.build(ctx, g44)
    cmap = pkg.PhotonMap.build(ctx, c44)
    return sc, s, g44, c44, gmap, cmap


def compare(fb, want):
    a = fb.view(np.uint8).astype(np.int32)
    b = want.view(np.uint8).astype(np.int32)
    diff = np.abs(a - b)
    return diff.max(), (diff != 0).mean()


@pytest.mark.parametrize("spp,depth", [(1, 1), (2, 3)])
def test_render_matches_oracle(pkg, ctx, orc, cornell, spp, depth):
    sc, s, g44, c44, gmap, cmap = cornell
    W, H = 96, 64
    cam = camera(pkg, W, H)
    fb = s.render(sc["lights"], gmap, cmap, cam, W, H, spp, depth)
    want, _ = oracle_render(orc, sc, g44, c44, cam, W, H, spp, depth)
    assert (fb[0] == 0).all() and (fb[1:] >> 24 == 0xFF).all()
    mx, frac = compare(fb, want)
    # stated tolerance: identical 8-bit image except isolated +-1 steps (none expected: the pipeline
    # uses the oracle's arithmetic operation for operation)
    assert mx <= 1 and frac < 1e-3, (mx, frac)


def test_render_matches_oracle_on_a_large_scene(pkg, ctx, orc):
    """a 6 000-triangle soup (above 4 096: Morton-built hierarchy, final-gather rays traced by the vote-driven warp
    traversal, k_fg<true>; the stored photons come from k_walk<true>): same image as the oracle's brute-force tracer"""
    from helpers import LIGHT, soup_scene

    sc = soup_scene(6000, 10, extent=6.0, size=1.8)
    sc["lights"] = np.zeros(2, LIGHT)  # on the camera's side of the soup
    sc["lights"][0] = ((3.0, 3.0, -6.0), (1, 1, 1), 40.0)
    sc["lights"][1] = ((-3.0, 8.0, -4.0), (1.0, 0.9, 0.8), 30.0)
    s = pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])
    g40, _, _, _ = s.emit_photons(sc["lights"], 20000, 8, False)
    c40, _, _, _ = s.emit_photons(sc["lights"], 5000, 8, True)
    g44, c44 = pkg.photons_to_maps(ctx, g40, c40)
    gmap, cmap = pkg.PhotonMap.build(ctx, g44), pkg.PhotonMap.build(ctx, c44)
    W, H = 64, 48
    cam = pkg.make_camera([3.0, 3.0, -9.0], [3.0, 3.0, 3.0], [0.0, 1.0, 0.0], 0.9, W, H)
    fb = s.render(sc["lights"], gmap, cmap, cam, W, H, 1, 2)
    want, _ = oracle_render(orc, sc, g44, c44, cam, W, H, 1, 2)
    mx, frac = compare(fb, want)
    assert (want[1:] != 0xFFFFFFFF).mean() > 0.5 and len(np.unique(want)) > 200  # the camera sees the soup, lit unevenly
    assert mx <= 1 and frac < 1e-3, (mx, frac)


def test_render_without_maps_and_tiles(pkg, ctx, orc, cornell):
    sc, s, g44, c44, gmap, cmap = cornell
    W, H = 64, 48
    cam = camera(pkg, W, H)
    fb = s.render(sc["lights"], None, None, cam, W, H, 1, 1, sky=(0.2, 0.4, 0.6))
    want, _ = oracle_render(orc, sc, None, None, cam, W, H, 1, 1, sky=(0.2, 0.4, 0.6))
    assert np.array_equal(fb, want)
    # tile sharding: three calls into one framebuffer equal one full call
    full = s.render(sc["lights"], gmap, cmap, cam, W, H, 1, 1)
    import ctypes
    tiled = np.zeros((H, W), np.uint32)
    lights = np.ascontiguousarray(sc["lights"], pkg.LIGHT)
    sky = np.ones(3, np.float32)
    for b, e in [(0, 1000), (1000, 1001), (1001, W * H)]:
        rc = pkg.lib().pm_render(ctx.handle, s.handle, lights.ctypes.data, len(lights), gmap.handle, cmap.handle,
                                 np.ascontiguousarray(cam).ctypes.data, W, H, 1, 1, sky.ctypes.data, b, e, tiled.ctypes.data)
        assert rc == 0
    assert np.array_equal(tiled, full)


# ---- parity at the BASELINE configuration sizes, with the image RMSE the north star asks for --------
def _rmse(a, b):
    d = a.astype(np.float64) - b.astype(np.float64)
    return float(np.sqrt((d * d).mean()))


# Stated bounds: the 8-bit images are identical except isolated +-1 steps; the RMSE of the linear
# colours (before make_rgba) is below 1e-4 absolute on images whose mean is about 0.03-0.3 — the device sums the
# k photon contributions of a gather in another order than the oracle's heap order, nothing else differs.
RMSE_BOUND = 1e-4


@pytest.fixture(scope="module")
def cornell_c1(pkg, ctx):
    """C1 as SURVEY §8(d) states it: cornell-box asset, 200 000 + 100 000 cast photons, max_depth 10, LCG parity RNG."""
    sc = load_scene("cornell_box")
    s = pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])
    g40, _, _, _ = s.emit_photons(sc["lights"], 200_000, 10, False)
    c40, _, _, _ = s.emit_photons(sc["lights"], 100_000, 10, True)
    g44, c44 = pkg.photons_to_maps(ctx, g40, c40)
    gmap = pkg.PhotonMap.build(ctx, g44)  # reorders g44 / c44 in place: the oracle gets the same tree arrays
    cmap = pkg.PhotonMap.build(ctx, c44)
    return sc, s, g44, c44, gmap, cmap


def test_c1_render_512_matches_oracle_with_rmse(pkg, ctx, orc, cornell_c1):
    sc, s, g44, c44, gmap, cmap = cornell_c1
    assert len(g44) > 300_000 and len(c44) > 1000
    W = H = 512
    cam = camera(pkg, W, H)
    fb, rgb = s.render(sc["lights"], gmap, cmap, cam, W, H, 1, 1, linear=True)
    want, want_rgb = oracle_render(orc, sc, g44, c44, cam, W, H, 1, 1)
    mx, frac = compare(fb, want)
    rmse = _rmse(rgb, want_rgb)
    print(f"C1 512x512: photons {len(g44)} + {len(c44)}, max 8-bit step {mx}, differing bytes {frac:.2e}, "
          f"linear RMSE {rmse:.3e} (image mean {want_rgb.mean():.4f})")
    assert mx <= 1 and frac < 1e-3, (mx, frac)
    assert rmse <= RMSE_BOUND, rmse


def test_c1_stress_spp24_depth30_crop_matches_oracle(pkg, ctx, orc, cornell_c1):
    """config.toml.example:16-17 values (samples_per_pixel 24, depth 30) on a 128x96 frame."""
    sc, s, g44, c44, gmap, cmap = cornell_c1
    W, H = 128, 96
    cam = camera(pkg, W, H)
    fb, rgb = s.render(sc["lights"], gmap, cmap, cam, W, H, 24, 30, linear=True)
    want, want_rgb = oracle_render(orc, sc, g44, c44, cam, W, H, 24, 30)
    mx, frac = compare(fb, want)
    rmse = _rmse(rgb, want_rgb)
    print(f"C1 stress 128x96 spp 24 depth 30: max 8-bit step {mx}, differing bytes {frac:.2e}, linear RMSE {rmse:.3e}")
    assert mx <= 1 and frac < 2e-3, (mx, frac)
    assert rmse <= RMSE_BOUND, rmse


def test_c3_reduced_1080p_range_matches_oracle(pkg, ctx, orc):
    """C3 at reduced photon count: 1 M cast photons + 100 k caustic, counter RNG seed 1, max_depth 8,
    1920x1080, one sixteenth of the frame (a band through the middle) through pixel_begin / pixel_end."""
    sc = load_scene("cornell_box")
    s = pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])
    g40, _, _, _ = s.emit_photons(sc["lights"], 1_000_000, 8, False, seed=1, rng_mode=pkg.RNG_COUNTER)
    c40, _, _, _ = s.emit_photons(sc["lights"], 100_000, 8, True, seed=1, rng_mode=pkg.RNG_COUNTER)
    g44, c44 = pkg.photons_to_maps(ctx, g40, c40)
    gmap = pkg.PhotonMap.build(ctx, g44)
    cmap = pkg.PhotonMap.build(ctx, c44)
    W, H = 1920, 1080
    cam = camera(pkg, W, H)
    n = W * H // 16
    b = (W * H - n) // 2
    fb, rgb = s.render(sc["lights"], gmap, cmap, cam, W, H, 1, 1, pixel_begin=b, pixel_end=b + n, linear=True)
    want, want_rgb = oracle_render(orc, sc, g44, c44, cam, W, H, 1, 1, begin=b, end=b + n)
    mx, frac = compare(fb, want)
    rmse = _rmse(rgb[b:b + n], want_rgb[b:b + n])
    print(f"C3 reduced: photons {len(g44)} + {len(c44)}, {n} pixels, max 8-bit step {mx}, differing bytes {frac:.2e}, "
          f"linear RMSE {rmse:.3e}")
    assert (rgb[:b] == 0).all() and (rgb[b + n:] == 0).all()
    assert mx <= 1 and frac < 1e-3, (mx, frac)
    assert rmse <= RMSE_BOUND, rmse

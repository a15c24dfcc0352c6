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

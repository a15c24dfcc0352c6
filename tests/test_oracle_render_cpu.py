"""CPU tests of the stage-3 oracle (oracle/orc_render.cpp): make_rgba table, framebuffer row quirk, sky
on miss, and hand-derived direct lighting. No reference golden image exists."""
import numpy as np

from helpers import LIGHT, MATERIAL, declare_render_oracle, load_scene, oracle_build, oracle_emit, oracle_render


def test_make_rgba_table(orc):
    declare_render_oracle(orc)
    f = orc.orc_make_rgba
    assert f(0.0, 0.0, 0.0) == 0xFF000000
    assert f(1.0, 1.0, 1.0) == 0xFFFFFFFF            # 256 clamps to 255
    assert f(0.5, 0.25, 0.999) == (0xFF << 24) | (255 << 16) | (64 << 8) | 128
    assert f(-1.0, float("nan"), 1e30) == (0xFF << 24) | (255 << 16) | (0 << 8) | 0
    assert f(0.00390625, 0.0039, 0.0) == 0xFF000001  # int(1.0) = 1, int(0.9984) = 0; linear, no gamma


def camera(pkg, W, H):
    return pkg.make_camera([80.0, 30.0, 0.0], [10.0, 20.0, 0.0], [0.0, 1.0, 0.0], 0.87, W, H)


def test_empty_scene_is_sky_and_row_quirk(pkg, orc):
    sc = load_scene("cornell_box")
    empty = dict(sc, verts=np.zeros((0, 3), np.float32), tris=np.zeros((0, 3), np.int32), tri_mesh=np.zeros(0, np.int32))
    W, H = 16, 8
    fb, rgb = oracle_render(orc, empty, None, None, camera(pkg, W, H), W, H, 2, 3, sky=(0.25, 0.5, 1.0))
    assert (fb[0] == 0).all()                                   # row 0 is never written (y = H - py >= 1)
    assert (fb[1:] == ((0xFF << 24) | (255 << 16) | (128 << 8) | 64)).all()
    assert np.allclose(rgb, [0.25, 0.5, 1.0])


def test_direct_light_value_on_floor(pkg, orc):
    """one white diffuse floor, one light straight above: direct = albedo * P * cos / d^2 * (diffuse / pi)"""
    v = np.float32([[-100, 0, -100], [100, 0, -100], [100, 0, 100], [-100, 0, 100]])
    t = np.int32([[0, 2, 1], [0, 3, 2]])
    mats = np.zeros(1, MATERIAL)
    mats[0] = ((0.5, 0.5, 0.5), 1.0, 0.0, 0.0, 1.0)
    lights = np.zeros(1, LIGHT)
    lights[0] = ((0, 10, 0), (1, 1, 1), 50.0)
    sc = dict(verts=v, tris=t, tri_mesh=np.zeros(2, np.int32), mats=mats, lights=lights)
    W = H = 9
    cam = pkg.make_camera([0.0, 5.0, 0.0], [0.0, 0.0, 0.0], [0.0, 0.0, 1.0], 1.5, W, H)  # narrow view straight down
    fb, rgb = oracle_render(orc, sc, None, None, cam, W, H, 1, 1)
    centre = rgb[(H // 2) * W + W // 2]
    want = 0.8 * 0.5 * 50.0 / 100.0 * (1.0 / np.pi)    # hit below the light: cos = 1, d = 10; specular spike adds 0
    assert np.allclose(centre, want, rtol=2e-3)


def test_render_uses_both_maps_and_is_deterministic(pkg, orc):
    sc = load_scene("cornell_box")
    g40, _, _ = oracle_emit(orc, sc, 3000, 8, False)
    c40, _, _ = oracle_emit(orc, sc, 3000, 8, True)
    from helpers import PHOTON44, vp
    g44, c44 = np.zeros(len(g40) + len(c40), PHOTON44), np.zeros(len(c40), PHOTON44)
    orc.orc_photons_to_maps(vp(g40), len(g40), vp(c40), len(c40), vp(g44), vp(c44))
    gt, _ = oracle_build(orc, g44)
    ct, _ = oracle_build(orc, c44)
    W, H = 24, 18
    cam = camera(pkg, W, H)
    fb0, rgb0 = oracle_render(orc, sc, None, None, cam, W, H, 1, 1)
    fb1, rgb1 = oracle_render(orc, sc, gt, ct, cam, W, H, 1, 1)
    fb2, rgb2 = oracle_render(orc, sc, gt, ct, cam, W, H, 1, 1)
    assert np.array_equal(fb1, fb2) and rgb1.tobytes() == rgb2.tobytes()
    assert (rgb1 >= rgb0 - 1e-6).all() and (rgb1 > rgb0 + 1e-4).any()     # photon terms only add light
    # tile sharding reproduces the full frame
    fb3 = np.zeros((H, W), np.uint32)
    for b, e in [(0, 100), (100, 250), (250, W * H)]:
        oracle_render(orc, sc, gt, ct, cam, W, H, 1, 1, begin=b, end=e, fb=fb3)
    assert np.array_equal(fb3, fb1)
    # depth 2 / spp 2 runs and differs
    fb4, _ = oracle_render(orc, sc, gt, ct, cam, W, H, 2, 2)
    assert not np.array_equal(fb4, fb1)

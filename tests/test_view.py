"""Photon viewer (≙ photonViewer, photon-viewer/src/hostCode.cu:53-75 +
photon-viewer/cuda/deviceCode.cu:10-49): projection pinned against the reference's own glm
(tests/golden/view_ref.npz, made by tests/golden/make_golden.py), splat compared with the CPU oracle."""
import ctypes
import os

import numpy as np
import pytest

from helpers import GOLDEN, PHOTON40, load_scene, vp

N_CASES = 3


def make_scene(pkg, ctx, sc):
    return pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])


def _cam(z, i):
    c = z[f"cam_{i}"]
    return c[0:3].copy(), c[3:6].copy(), c[6:9].copy(), float(c[9]), int(c[10]), int(c[11])


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(GOLDEN, "view_ref.npz"))


def _declare(orc):
    f, i64, p, ci = ctypes.c_float, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int
    orc.orc_view_matrix.argtypes = [p, p, p, f, ci, ci, p]
    orc.orc_view_project.argtypes = [p, p, p, f, ci, ci, p, i64, p]
    orc.orc_view_photons.argtypes = [p, ci, p, ci, p, i64, p, p, p, f, ci, ci, p]
    return orc


@pytest.mark.parametrize("i", range(N_CASES))
def test_oracle_projection_matches_glm(orc, golden, i):
    _declare(orc)
    frm, at, up, fovy, W, H = _cam(golden, i)
    m = np.zeros(16, np.float32)
    orc.orc_view_matrix(vp(frm), vp(at), vp(up), fovy, W, H, vp(m))
    assert m.tobytes() == golden[f"matrix_{i}"].tobytes()          # bit-exact P * V
    pos = np.ascontiguousarray(golden[f"pos_{i}"])
    pix = np.zeros((len(pos), 2), np.int32)
    orc.orc_view_project(vp(frm), vp(at), vp(up), fovy, W, H, vp(pos), len(pos), vp(pix))
    want = golden[f"pix_{i}"]
    assert np.array_equal(pix, want)
    inside = (want[:, 0] >= 0) & (want[:, 0] < W) & (want[:, 1] >= 0) & (want[:, 1] < H)
    assert inside.sum() > 100                                       # the case really lands photons on screen


@pytest.mark.parametrize("i", range(N_CASES))
def test_library_matrix_matches_glm(pkg, golden, i):
    """pm_view_matrix is host arithmetic: it runs without a GPU."""
    frm, at, up, fovy, W, H = _cam(golden, i)
    m = np.zeros(16, np.float32)
    pkg.lib().pm_view_matrix(vp(frm), vp(at), vp(up), fovy, W, H, vp(m))
    assert m.tobytes() == golden[f"matrix_{i}"].tobytes()


def _oracle_view(orc, sc, photons, frm, at, up, fovy, W, H):
    _declare(orc)
    fb = np.zeros((H, W), np.uint32)
    orc.orc_view_photons(vp(sc["verts"]), len(sc["verts"]), vp(sc["tris"]), len(sc["tris"]), vp(photons), len(photons),
                         vp(frm), vp(at), vp(up), fovy, W, H, vp(fb))
    return fb


@pytest.mark.gpu
@pytest.mark.parametrize("scene_name,caustics", [("cornell_box", False), ("cornell_box", True), ("sphere", False)])
def test_view_matches_oracle(pkg, ctx, orc, scene_name, caustics):
    sc = load_scene(scene_name)
    s = make_scene(pkg, ctx, sc)
    photons, cnt, _, _ = s.emit_photons(sc["lights"], 40000 if scene_name == "cornell_box" else 4000, 8, caustics)
    photons = photons[:cnt].copy()
    lo, hi = sc["verts"].min(0), sc["verts"].max(0)
    ctr, ext = (lo + hi) / 2, float((hi - lo).max())
    frm = (ctr + np.array([1.6 * ext, 0.3 * ext, 0.2 * ext])).astype(np.float32)
    at, up, fovy, W, H = ctr.astype(np.float32), np.array([0, 1, 0], np.float32), 0.87, 256, 192
    got = s.view_photons(photons, frm, at, up, fovy, W, H)
    want = _oracle_view(orc, sc, photons, frm, at, up, fovy, W, H)
    lit = (want != 0xFF000000).sum()
    if scene_name == "cornell_box" and not caustics:
        assert lit > 2000                      # the image is not empty
    # norm3d (device libm) vs the oracle's rounded double length: tmax may differ by one ulp for a photon whose
    # own surface hit sits exactly there; allow a handful of flipped pixels, everything else is bit-exact
    assert (got != want).sum() <= max(2, lit // 2000)


@pytest.mark.gpu
def test_view_edge_cases(pkg, ctx):
    sc = load_scene("cornell_box")
    s = make_scene(pkg, ctx, sc)
    frm, at, up = np.array([80, 30, 0], np.float32), np.array([10, 20, 0], np.float32), np.array([0, 1, 0], np.float32)
    empty = s.view_photons(np.zeros(0, PHOTON40), frm, at, up, 0.87, 64, 48)
    assert empty.shape == (48, 64) and (empty == 0xFF000000).all()
    # photons behind the camera or outside the frustum never write; two photons at one place: the later one wins
    ph = np.zeros(4, PHOTON40)
    ph["pos"] = [[200, 30, 0], [10, 20, 5000], [79, 29.9, 0], [79, 29.9, 0]]
    ph["color"] = [[1, 1, 1], [1, 1, 1], [1, 0, 0], [0, 0.5, 0]]
    fb = s.view_photons(ph, frm, at, up, 0.87, 64, 48)
    lit = fb[fb != 0xFF000000]
    assert len(lit) == 1 and lit[0] == (0xFF000000 | (128 << 8))
    # device-resident photons
    import torch
    d = torch.from_numpy(ph.view(np.uint8)).cuda()
    fb2 = np.empty((48, 64), np.uint32)
    pkg._check(pkg.lib().pm_view_photons(ctx.handle, s.handle, d.data_ptr(), 4, frm.ctypes.data, at.ctypes.data,
                                         up.ctypes.data, 0.87, 64, 48, fb2.ctypes.data))
    assert np.array_equal(fb, fb2)

"""Procedural scenes of the BASELINE configs that have no asset in the reference: the C4 room — a closed
box of Cornell proportions (40^3 units, walls facing inwards) around a displaced sphere tessellated to about one million
triangles, 7 diffuse-dominant materials, 2 point lights. Deterministic formulae, no files; used by bench.py, the tests
and tools/c4_run.py. numpy only (needs no GPU library)."""
from __future__ import annotations

import numpy as np

MATERIAL = np.dtype(
    [("albedo", "<f4", 3), ("diffuse", "<f4"), ("specular", "<f4"), ("transmission", "<f4"), ("refraction_idx", "<f4")]
)
LIGHT = np.dtype([("pos", "<f4", 3), ("rgb", "<f4", 3), ("power", "<f8")])


def displaced_sphere(nu: int, nv: int, radius: float, centre):
    """Sphere of 2*nu*nv triangles whose radius is modulated by a sum of four sine products."""
    u = np.linspace(0, np.pi, nu + 1, dtype=np.float64)[:, None]
    v = np.linspace(0, 2 * np.pi, nv + 1, dtype=np.float64)[None, :]
    r = radius * (1 + 0.08 * np.sin(7 * u) * np.sin(5 * v) + 0.04 * np.sin(23 * u + 1.3) * np.sin(17 * v) +
                  0.02 * np.sin(61 * u) * np.sin(53 * v + 0.7) + 0.01 * np.sin(131 * u + 2.1) * np.sin(113 * v))
    p = np.stack([r * np.sin(u) * np.cos(v), r * np.cos(u) + 0 * v, r * np.sin(u) * np.sin(v)], -1) + np.asarray(centre)
    idx = np.arange((nu + 1) * (nv + 1)).reshape(nu + 1, nv + 1)
    a, b, c, d = idx[:-1, :-1], idx[1:, :-1], idx[1:, 1:], idx[:-1, 1:]
    tris = np.concatenate([np.stack([a, b, c], -1).reshape(-1, 3), np.stack([a, c, d], -1).reshape(-1, 3)])
    return p.reshape(-1, 3).astype(np.float32), tris.astype(np.int32)


def box(lo, hi):
    """12 triangles wound so that the geometric normals point INTO the room: the reference scatters about the
    unflipped normal (common/cuda/helpers.h:45-47,76-85), a room whose walls face outwards loses every photon."""
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    c = np.array([[x, y, z] for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])], np.float32)
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    tris = np.array([t for q in quads for t in ((q[0], q[2], q[1]), (q[0], q[3], q[2]))], np.int32)
    return c, tris


def c4_room(nu: int = 500, nv: int = 1000):
    """The C4 scene: 2*nu*nv + 12 triangles (default 1 000 012). Returns a dict like tests' load_scene():
    verts, tris, tri_mesh, mats (MATERIAL), lights (LIGHT), camera (look_from, look_at, look_up, fovy)."""
    sv, st = displaced_sphere(nu, nv, 12.0, (0.0, 14.0, 0.0))
    bv, bt = box((-20, 0, -20), (20, 40, 20))
    verts = np.concatenate([sv, bv])
    tris = np.concatenate([st, bt + len(sv)])
    tri_mesh = np.concatenate([np.zeros(len(st), np.int32), 1 + (np.arange(len(bt)) // 2).astype(np.int32)])
    mats = np.zeros(7, MATERIAL)
    mats[0] = ((0.8, 0.7, 0.6), 0.7, 0.2, 0.0, 1.0)
    for i in range(1, 7):
        mats[i] = ((0.75, 0.75, 0.75) if i > 2 else ((0.8, 0.2, 0.2) if i == 1 else (0.2, 0.8, 0.2)), 0.9, 0.0, 0.0, 1.0)
    lights = np.zeros(2, LIGHT)
    lights[0] = ((-12.0, 36.0, -8.0), (1, 1, 1), 10.0)
    lights[1] = ((13.0, 35.0, 11.0), (1.0, 0.9, 0.8), 10.0)
    return dict(verts=np.ascontiguousarray(verts, np.float32), tris=np.ascontiguousarray(tris, np.int32),
                tri_mesh=np.ascontiguousarray(tri_mesh, np.int32), mats=mats, lights=lights,
                camera=((0.0, 20.0, -19.0), (0.0, 14.0, 0.0), (0.0, 1.0, 0.0), 0.87))

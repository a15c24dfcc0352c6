"""Shared test helpers: oracle loading with typed signatures, synthetic clouds, comparison utilities."""
import ctypes
import os
import subprocess
from ctypes import POINTER, c_float, c_int, c_int32, c_int64, c_uint8, c_uint32, c_uint64, c_void_p

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

PHOTON44 = np.dtype(
    [("pos", "<f4", 3), ("dir", "<f4", 3), ("color", "<f4", 3), ("power", "<f4"), ("qn", "u1", 3), ("split_dim", "u1")]
)
PHOTON40 = np.dtype([("pos", "<f4", 3), ("dir", "<f4", 3), ("power", "<i4"), ("color", "<f4", 3)])


def load_oracle():
    path = os.path.join(ROOT, "oracle", "libpm_oracle.so")
    if not os.path.exists(path):
        env = {k: v for k, v in os.environ.items() if k not in ("CC", "CXX")}
        subprocess.run(["make"], cwd=os.path.join(ROOT, "oracle"), check=True, env=env)
    L = ctypes.CDLL(path)
    L.orc_subtree_size.restype = c_int64
    L.orc_subtree_size.argtypes = [c_int64, c_int64]
    L.orc_kd_build.argtypes = [c_void_p, c_int64, c_void_p]
    L.orc_kd_check.argtypes = [c_void_p, c_int]
    L.orc_knn_traverse.restype = c_float
    L.orc_knn_traverse.argtypes = [c_void_p, c_int, c_void_p, c_int, c_float, c_void_p, POINTER(c_int64)]
    L.orc_knn_brute.restype = c_float
    L.orc_knn_brute.argtypes = [c_void_p, c_int, c_void_p, c_int, c_float, c_void_p, POINTER(c_int)]
    L.orc_knn_batch.argtypes = [c_void_p, c_int, c_void_p, c_int64, c_int, c_float, c_void_p, c_void_p, c_void_p, c_int]
    L.orc_gather_batch.argtypes = [c_void_p, c_int, c_void_p, c_void_p, c_int64, c_int, c_float, c_void_p, c_int]
    declare_scene_oracle(L)
    return L


def vp(a):
    return a.ctypes.data_as(c_void_p) if a is not None else None


def unique_cloud(n, seed, scale=2.0**-8):
    """Photons with per-axis UNIQUE coordinates (a permutation of 0..n-1 times a power of two), the
    construction cudaKDTree/testing/testBuildersSameResult.cu:96-115 uses to make the tree unique."""
    rng = np.random.default_rng(seed)
    a = np.zeros(n, PHOTON44)
    for d in range(3):
        a["pos"][:, d] = rng.permutation(n).astype(np.float32) * np.float32(scale)
    a["dir"] = rng.random((n, 3), dtype=np.float32)
    a["color"] = rng.random((n, 3), dtype=np.float32)
    a["power"] = np.where(rng.random(n) < 0.8, 1.0, 0.5).astype(np.float32)
    a["qn"] = rng.integers(0, 255, (n, 3), dtype=np.uint8)
    a["split_dim"] = 255
    return a


def dup_cloud(n, seed, levels=6):
    """Photons on a coarse integer lattice: many coordinate ties (tie-break = input index)."""
    rng = np.random.default_rng(seed)
    a = np.zeros(n, PHOTON44)
    a["pos"] = rng.integers(0, levels, (n, 3)).astype(np.float32)
    a["color"] = rng.random((n, 3), dtype=np.float32)
    a["power"] = 1.0
    a["dir"][:, 0] = np.arange(n, dtype=np.float32)  # carries the input index through the build
    return a


def oracle_build(orc, photons):
    t = photons.copy()
    b = np.zeros(6, np.float32)
    orc.orc_kd_build(vp(t), len(t), vp(b))
    return t, b


def oracle_knn(orc, tree, queries, k, cutoff):
    nq = len(queries)
    ids = np.empty((nq, k), np.int32)
    d2 = np.empty((nq, k), np.float32)
    r2 = np.empty(nq, np.float32)
    q = np.ascontiguousarray(queries, np.float32)
    orc.orc_knn_batch(vp(tree), len(tree), vp(q), nq, k, cutoff, vp(ids), vp(d2), vp(r2), 0)
    return ids, d2, r2


def oracle_brute(orc, tree, query, k, cutoff):
    out = np.empty(k, np.uint64)
    tie = c_int(0)
    q = np.ascontiguousarray(query, np.float32)
    r2 = orc.orc_knn_brute(vp(tree), len(tree), vp(q), k, cutoff, vp(out), ctypes.byref(tie))
    return out, r2, tie.value


def entries(ids, d2):
    """(dist² bits << 32 | id) entries sorted per row — the canonical comparison form."""
    e = (d2.view(np.uint32).astype(np.uint64) << np.uint64(32)) | ids.view(np.uint32).astype(np.uint64)
    return np.sort(e, axis=1)


# ---- scenes (stage 1 / 3) ---------------------------------------------------------------------
MATERIAL = np.dtype(
    [("albedo", "<f4", 3), ("diffuse", "<f4"), ("specular", "<f4"), ("transmission", "<f4"), ("refraction_idx", "<f4")]
)
LIGHT = np.dtype([("pos", "<f4", 3), ("rgb", "<f4", 3), ("power", "<f8")])
RAY = np.dtype([("org", "<f4", 3), ("tmin", "<f4"), ("dir", "<f4", 3), ("tmax", "<f4")])
HIT = np.dtype([("prim", "<i4"), ("t", "<f4")])


def load_scene(name):
    """Committed fixture extracted from the reference's assets by tests/golden/make_scenes.py."""
    z = np.load(os.path.join(GOLDEN, f"scene_{name}.npz"))
    return dict(verts=np.ascontiguousarray(z["verts"]), tris=np.ascontiguousarray(z["tris"]),
                tri_mesh=np.ascontiguousarray(z["tri_mesh"]), mats=z["mats"].view(MATERIAL).reshape(-1).copy(),
                lights=z["lights"].view(LIGHT).reshape(-1).copy(), names=list(z["names"]))


def soup_scene(nt, seed, extent=10.0, size=1.5):
    """Random triangle soup with 4 materials (diffuse / mirror / glass / mixed)."""
    rng = np.random.default_rng(seed)
    c = rng.random((nt, 1, 3), dtype=np.float32) * np.float32(extent)
    verts = (c + (rng.random((nt, 3, 3), dtype=np.float32) - np.float32(0.5)) * np.float32(size)).reshape(-1, 3)
    tris = np.arange(3 * nt, dtype=np.int32).reshape(nt, 3)
    tri_mesh = rng.integers(0, 4, nt).astype(np.int32)
    mats = np.zeros(4, MATERIAL)
    mats[0] = ((0.7, 0.6, 0.5), 0.9, 0.0, 0.0, 1.0)
    mats[1] = ((1.0, 1.0, 1.0), 0.0, 1.0, 0.0, 1.0)
    mats[2] = ((0.8, 0.8, 1.0), 0.0, 0.2, 0.8, 1.5)
    mats[3] = ((0.9, 0.9, 0.9), 0.5, 0.2, 0.2, 1.3)
    lights = np.zeros(2, LIGHT)
    lights[0] = ((extent * 0.5, extent * 0.5, extent * 0.5), (1, 1, 1), 10.0)
    lights[1] = ((extent * 0.25, extent * 0.75, extent * 0.3), (1.0, 0.9, 0.8), 5.0)
    return dict(verts=np.ascontiguousarray(verts, np.float32), tris=tris, tri_mesh=tri_mesh, mats=mats, lights=lights)


def random_rays(n, seed, lo, hi, tmin=1e-3, tmax=1e30):
    rng = np.random.default_rng(seed)
    r = np.zeros(n, RAY)
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    r["org"] = lo + rng.random((n, 3), dtype=np.float32) * (hi - lo)
    d = rng.standard_normal((n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    # a share of axis-parallel directions (zero components: 0 * inf in the slab test)
    ax = rng.integers(0, 8, n)
    for k in range(3):
        sel = ax == k
        d[sel] = 0
        d[sel, k] = np.where(rng.random(sel.sum()) < 0.5, -1, 1)
    r["dir"] = d
    r["tmin"], r["tmax"] = tmin, tmax
    return r


def declare_scene_oracle(L):
    L.orc_trace_closest.argtypes = [c_void_p, c_int, c_void_p, c_int, c_void_p, c_int64, c_void_p]
    L.orc_trace_occluded.argtypes = [c_void_p, c_int, c_void_p, c_int, c_void_p, c_int64, c_void_p]
    L.orc_light_paths.restype = c_int64
    L.orc_light_paths.argtypes = [c_void_p, c_int, c_int64, c_int]
    L.orc_emit_photons.restype = c_int64
    L.orc_emit_photons.argtypes = [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int64,
                                   c_int, c_int, c_uint64, c_int, c_int64, c_int64, c_void_p, c_int64, POINTER(c_int64),
                                   POINTER(c_int64)]
    L.orc_photons_to_maps.argtypes = [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p]
    return L


def oracle_trace(orc, sc, rays):
    hits = np.empty(len(rays), HIT)
    orc.orc_trace_closest(vp(sc["verts"]), len(sc["verts"]), vp(sc["tris"]), len(sc["tris"]), vp(rays), len(rays), vp(hits))
    return hits


def oracle_occluded(orc, sc, rays):
    occ = np.empty(len(rays), np.uint8)
    orc.orc_trace_occluded(vp(sc["verts"]), len(sc["verts"]), vp(sc["tris"]), len(sc["tris"]), vp(rays), len(rays), vp(occ))
    return occ


LIGHT_SHAPE = np.dtype([("type", "<i4"), ("normal", "<f4", 3), ("side_length", "<f4")])


def oracle_emit(orc, sc, casted, max_depth, caustics, seed=0, rng_mode=0, first=0, stride=1, lights=None, shapes=None):
    lights = sc["lights"] if lights is None else lights
    cap = max(1, int(casted) * max(1, max_depth))
    out = np.zeros(cap, PHOTON40)
    npaths, nseg = c_int64(0), c_int64(0)
    if shapes is not None:
        orc.orc_emit_photons_shaped.restype = c_int64
        orc.orc_emit_photons_shaped.argtypes = [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                                c_int, c_int64, c_int, c_int, c_uint64, c_int, c_int64, c_int64, c_void_p,
                                                c_int64, POINTER(c_int64), POINTER(c_int64)]
        shapes = np.ascontiguousarray(shapes, LIGHT_SHAPE)
        n = orc.orc_emit_photons_shaped(vp(sc["verts"]), len(sc["verts"]), vp(sc["tris"]), len(sc["tris"]), vp(sc["tri_mesh"]),
                                        vp(sc["mats"]), len(sc["mats"]), vp(lights), vp(shapes), len(lights), casted, max_depth,
                                        1 if caustics else 0, seed, rng_mode, first, stride, vp(out), cap,
                                        ctypes.byref(npaths), ctypes.byref(nseg))
        return out[:n].copy(), npaths.value, nseg.value
    n = orc.orc_emit_photons(vp(sc["verts"]), len(sc["verts"]), vp(sc["tris"]), len(sc["tris"]), vp(sc["tri_mesh"]),
                             vp(sc["mats"]), len(sc["mats"]), vp(lights), len(lights), casted, max_depth,
                             1 if caustics else 0, seed, rng_mode, first, stride, vp(out), cap, ctypes.byref(npaths),
                             ctypes.byref(nseg))
    return out[:n].copy(), npaths.value, nseg.value


# ---- stage 3 (camera pass) --------------------------------------------------------------------
CAMERA = np.dtype([("pos", "<f4", 3), ("dir_00", "<f4", 3), ("dir_du", "<f4", 3), ("dir_dv", "<f4", 3)])


def declare_render_oracle(L):
    L.orc_make_rgba.restype = c_uint32
    L.orc_make_rgba.argtypes = [c_float, c_float, c_float]
    L.orc_render.argtypes = [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_int,
                             c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int64, c_int64, c_void_p,
                             c_void_p]
    return L


def oracle_render(orc, sc, gmap, cmap, cam, W, H, spp, depth, sky=(1, 1, 1), begin=0, end=None, fb=None):
    """gmap / cmap: tree-ordered PHOTON44 arrays (or None)."""
    declare_render_oracle(orc)
    end = W * H if end is None else end
    fb = np.zeros((H, W), np.uint32) if fb is None else fb
    rgb = np.zeros((W * H, 3), np.float32)
    sky = np.asarray(sky, np.float32)
    cam = np.ascontiguousarray(cam, CAMERA)
    orc.orc_render(vp(sc["verts"]), len(sc["verts"]), vp(sc["tris"]), len(sc["tris"]), vp(sc["tri_mesh"]), vp(sc["mats"]),
                   len(sc["mats"]), vp(sc["lights"]), len(sc["lights"]),
                   vp(gmap) if gmap is not None and len(gmap) else None, 0 if gmap is None else len(gmap),
                   vp(cmap) if cmap is not None and len(cmap) else None, 0 if cmap is None else len(cmap),
                   vp(cam), W, H, spp, depth, vp(sky), begin, end, vp(fb), vp(rgb))
    return fb, rgb


# ---- C++ hosts (photon-mapping_b200/host) -------------------------------------------------------
HOST_DIR = os.path.join(ROOT, "photon-mapping_b200", "host")


def write_glb(path, meshes, node_transforms=None):
    """Minimal binary glTF writer for tests. meshes: list of (material_name, verts[nv,3] f32, tris[nt,3]);
    node_transforms: optional list of dicts (translation / rotation / scale) per mesh node."""
    import json
    import struct

    bin_parts, views, accessors, gl_meshes, nodes, materials = [], [], [], [], [], []
    off = 0
    for i, (name, v, t) in enumerate(meshes):
        v = np.ascontiguousarray(v, np.float32)
        idx = np.ascontiguousarray(t, np.uint32).reshape(-1)
        for data, target in ((v.tobytes(), 34962), (idx.tobytes(), 34963)):
            pad = (-len(data)) % 4
            views.append({"buffer": 0, "byteOffset": off, "byteLength": len(data), "target": target})
            bin_parts.append(data + b"\0" * pad)
            off += len(data) + pad
        accessors.append({"bufferView": 2 * i, "componentType": 5126, "count": len(v), "type": "VEC3",
                          "min": v.min(0).tolist(), "max": v.max(0).tolist()})
        accessors.append({"bufferView": 2 * i + 1, "componentType": 5125, "count": len(idx), "type": "SCALAR"})
        materials.append({"name": name})
        gl_meshes.append({"primitives": [{"attributes": {"POSITION": 2 * i}, "indices": 2 * i + 1, "material": i}]})
        node = {"mesh": i, "name": f"node{i}"}
        if node_transforms and node_transforms[i]:
            node.update(node_transforms[i])
        nodes.append(node)
    root = {"name": "root", "children": list(range(len(nodes)))}
    nodes.append(root)
    j = {"asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [len(nodes) - 1]}], "nodes": nodes,
         "meshes": gl_meshes, "materials": materials, "accessors": accessors, "bufferViews": views,
         "buffers": [{"byteLength": off}]}
    js = json.dumps(j).encode()
    js += b" " * ((-len(js)) % 4)
    blob = b"".join(bin_parts)
    total = 12 + 8 + len(js) + 8 + len(blob)
    with open(path, "wb") as f:
        f.write(struct.pack("<III", 0x46546C67, 2, total))
        f.write(struct.pack("<II", len(js), 0x4E4F534A) + js)
        f.write(struct.pack("<II", len(blob), 0x004E4942) + blob)


def write_scene_files(dirpath, sc, name="scene", node_transforms=None):
    """Writes <dir>/<name>.glb, <name>.mtl and lights.txt for a fixture scene dict (load_scene); returns the glb path."""
    meshes = []
    for mi in range(len(sc["mats"])):
        sel = sc["tri_mesh"] == mi
        tris = sc["tris"][sel]
        used, inv = np.unique(tris.reshape(-1), return_inverse=True)
        meshes.append((f"mat{mi}", sc["verts"][used], inv.reshape(-1, 3)))
    glb = os.path.join(dirpath, name + ".glb")
    write_glb(glb, meshes, node_transforms)
    with open(os.path.join(dirpath, name + ".mtl"), "w") as f:
        f.write("# material_name albedo_r albedo_g albedo_b diffuse specular transmission ior\n\n")
        for mi, m in enumerate(sc["mats"]):
            f.write(f"mat{mi} {m['albedo'][0]:.9g} {m['albedo'][1]:.9g} {m['albedo'][2]:.9g} {m['diffuse']:.9g} "
                    f"{m['specular']:.9g} {m['transmission']:.9g} {m['refraction_idx']:.9g}\n")
    with open(os.path.join(dirpath, "lights.txt"), "w") as f:
        f.write("# posx posy posz colourr colourg colourb power\n")
        for l in sc["lights"]:
            f.write(" ".join(f"{x:.9g}" for x in (*l["pos"], *l["rgb"], l["power"])) + "\n")
    return glb


def write_config(path, model, photons, caustics, out_png, W=64, H=48, spp=1, depth=1, max_depth=6, casted=(20000, 10000),
                 camera=((80.0, 30.0, 0.0), (10.0, 20.0, 0.0), (0.0, 1.0, 0.0), 0.87)):
    frm, at, up, fovy = camera
    with open(path, "w") as f:
        f.write(f"""[camera]
look_from = [{frm[0]}, {frm[1]}, {frm[2]}]
look_at = [{at[0]}, {at[1]}, {at[2]}]
look_up = [{up[0]}, {up[1]}, {up[2]}]
fovy = {fovy}

[data]
photons_file = "{photons}"   # written by photonMapping
caustics_photons_file = "{caustics}"
model_path = "{model}"

[ray-tracer]
sky_colour = [1.0, 1.0, 1.0]
output_filename = "{out_png}"
fb_size = [{W}, {H}]
samples_per_pixel = {spp}
depth = {depth}

[photon-mapper]
max_depth = {max_depth}
casted_diffuse_photons = {casted[0]:_}
casted_caustics_photons = {casted[1]:_}

[photon-viewer]
output_filename = "{out_png}.viewer.png"
caustics_output_filename = "{out_png}.caustics-viewer.png"
fb_size = [{W}, {H}]
""")


def read_png_rgba8(path):
    """Decoder for the PNGs the hosts write (8-bit RGBA, filter 0)."""
    import struct
    import zlib

    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    off, idat, W, H = 8, b"", 0, 0
    while off < len(b):
        n, typ = struct.unpack(">I4s", b[off:off + 8])
        data = b[off + 8:off + 8 + n]
        assert zlib.crc32(typ + data) == struct.unpack(">I", b[off + 8 + n:off + 12 + n])[0]
        if typ == b"IHDR":
            W, H, depth, ctype = struct.unpack(">IIBB", data[:10])
            assert depth == 8 and ctype == 6
        elif typ == b"IDAT":
            idat += data
        off += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(H, 1 + 4 * W)
    assert (raw[:, 0] == 0).all()
    return np.ascontiguousarray(raw[:, 1:]).view("<u4").reshape(H, W)

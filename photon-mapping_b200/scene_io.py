"""Scene loading for the Python host: mirrors ``assets::import_scene`` (common/src/assetImporter.cxx:16-215)
for binary glTF (.glb) files without assimp.

What the reference does and this reproduces:
  * every mesh primitive of every node becomes one ``Mesh`` whose vertices are the node-transformed
    positions, de-duplicated per mesh by exact float equality in first-seen order
    (assetImporter.cxx:54-78);
  * ``lights.txt`` beside the model: ``posx posy posz r g b power`` per line, '#' comments
    (assetImporter.cxx:98-137); a missing file raises;
  * ``<model>.mtl``: ``name r g b diffuse specular transmission ior``; meshes whose material name is
    not listed get the white diffuse default (assetImporter.cxx:139-214); a missing file only warns.
Deviation (SURVEY §8a quirk viii): paths are used as given; the reference rewrites '/' to '\\'.
"""
from __future__ import annotations

import json
import os
import struct
import sys

import numpy as np

MATERIAL = np.dtype(
    [("albedo", "<f4", 3), ("diffuse", "<f4"), ("specular", "<f4"), ("transmission", "<f4"), ("refraction_idx", "<f4")]
)
LIGHT = np.dtype([("pos", "<f4", 3), ("rgb", "<f4", 3), ("power", "<f8")])
LIGHT_SHAPE = np.dtype([("type", "<i4"), ("normal", "<f4", 3), ("side_length", "<f4")])

_COMP = {5120: "i1", 5121: "u1", 5122: "<i2", 5123: "<u2", 5125: "<u4", 5126: "<f4"}
_NCOMP = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


def _accessor(j, blob, idx):
    acc = j["accessors"][idx]
    bv = j["bufferViews"][acc["bufferView"]]
    dt = np.dtype(_COMP[acc["componentType"]])
    nc = _NCOMP[acc["type"]]
    off = bv.get("byteOffset", 0) + acc.get("byteOffset", 0)
    stride = bv.get("byteStride", 0) or dt.itemsize * nc
    cnt = acc["count"]
    if stride == dt.itemsize * nc:
        return np.frombuffer(blob, dt, cnt * nc, off).reshape(cnt, nc)
    out = np.empty((cnt, nc), dt)
    for i in range(cnt):
        out[i] = np.frombuffer(blob, dt, nc, off + i * stride)
    return out


def _node_matrix(node):
    """Local transform of a glTF node as float32 4x4 (T * R * S or the explicit matrix)."""
    f32 = np.float32
    if "matrix" in node:
        return np.array(node["matrix"], f32).reshape(4, 4).T.copy()
    t = np.array(node.get("translation", [0, 0, 0]), f32)
    q = np.array(node.get("rotation", [0, 0, 0, 1]), f32)
    s = np.array(node.get("scale", [1, 1, 1]), f32)
    x, y, z, w = (f32(v) for v in q)
    r = np.array(
        [
            [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
            [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
            [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)],
        ],
        f32,
    )
    m = np.eye(4, dtype=f32)
    m[:3, :3] = r * s[None, :]
    m[:3, 3] = t
    return m


def load_glb_meshes(path: str):
    """Returns a list of (material_name, vertices[nv,3] f32, indices[nt,3] i32) in BFS node order."""
    blob = open(path, "rb").read()
    magic, _ver, _length = struct.unpack_from("<III", blob, 0)
    if magic != 0x46546C67:
        raise ValueError(f"{path}: not a binary glTF file")
    off = 12
    j, bin_chunk = None, b""
    while off < len(blob):
        clen, ctype = struct.unpack_from("<II", blob, off)
        data = blob[off + 8 : off + 8 + clen]
        if ctype == 0x4E4F534A:
            j = json.loads(data)
        elif ctype == 0x004E4942:
            bin_chunk = data
        off += 8 + clen
    scene = j["scenes"][j.get("scene", 0)]
    queue = [(n, np.eye(4, dtype=np.float32)) for n in scene["nodes"]]
    meshes = []
    while queue:  # breadth first, like assetImporter.cxx:39-47
        ni, parent = queue.pop(0)
        node = j["nodes"][ni]
        m = (parent @ _node_matrix(node)).astype(np.float32)
        for c in node.get("children", []):
            queue.append((c, m))
        if "mesh" not in node:
            continue
        for prim in j["meshes"][node["mesh"]]["primitives"]:
            if prim.get("mode", 4) != 4:
                continue  # aiProcess_SortByPType: only triangles reach the loop that asserts 3 indices
            pos = _accessor(j, bin_chunk, prim["attributes"]["POSITION"]).astype(np.float32)
            if "indices" in prim:
                idx = _accessor(j, bin_chunk, prim["indices"]).astype(np.int64).reshape(-1)
            else:
                idx = np.arange(len(pos), dtype=np.int64)
            world = (pos @ m[:3, :3].T + m[:3, 3]).astype(np.float32)
            # per-mesh de-duplication by exact equality, first-seen order over the face stream
            face_verts = world[idx]
            key = np.ascontiguousarray(face_verts).view([("", "<u4")] * 3).reshape(-1)
            _, first, inv = np.unique(key, return_index=True, return_inverse=True)
            order = np.argsort(first, kind="stable")
            rank = np.empty_like(order)
            rank[order] = np.arange(len(order))
            verts = face_verts[first[order]]
            tris = rank[inv].reshape(-1, 3).astype(np.int32)
            mat = prim.get("material")
            name = j["materials"][mat].get("name", "") if mat is not None else ""
            meshes.append((name, verts, tris))
    return meshes


def read_light_shapes(path: str) -> np.ndarray:
    """Emitter shapes of lights.txt: lines with four extra values "nx ny nz side" are square emitters (type 1)."""
    rows = [ln.split() for ln in open(path) if ln.strip() and ln[0] != "#"]
    shapes = np.zeros(len(rows), LIGHT_SHAPE)
    for i, tok in enumerate(rows):
        if len(tok) >= 11:
            shapes[i] = (1, tuple(float(t) for t in tok[7:10]), float(tok[10]))
    return shapes


def read_lights(path: str) -> np.ndarray:
    if not os.path.exists(path):
        raise RuntimeError("Unable to open file: " + path)
    out = []
    for line in open(path):
        line = line.rstrip("\n")
        if not line or line[0] == "#":
            continue
        tok = line.split()
        try:
            vals = [float(t) for t in tok[:7]]
            if len(vals) < 7:
                raise ValueError
        except ValueError:
            raise RuntimeError("Invalid light source data format")
        out.append(vals)
    lights = np.zeros(len(out), LIGHT)
    for i, v in enumerate(out):
        lights[i]["pos"], lights[i]["rgb"], lights[i]["power"] = v[0:3], v[3:6], v[6]
    return lights


def read_materials(path: str) -> dict:
    mats = {}
    if not os.path.exists(path):
        print(f"Error: Unable to open file {path}", file=sys.stderr)
        return mats
    for line in open(path):
        line = line.rstrip("\n")
        if not line or line[0] == "#":
            continue
        tok = line.split()
        try:
            name, vals = tok[0], [float(t) for t in tok[1:8]]
            if len(vals) < 7:
                raise ValueError
        except (ValueError, IndexError):
            print(f"Warning: Invalid line format: {line}", file=sys.stderr)
            continue
        mats[name] = vals
    return mats


def import_scene(model_path: str):
    """≙ assets::import_scene. Returns dict(verts, tris, tri_mesh, mats, lights, names) with all meshes
    concatenated (vertex indices rebased), ready for ``Scene(ctx, verts, tris, tri_mesh, mats)``."""
    meshes = load_glb_meshes(model_path)
    lights = read_lights(os.path.join(os.path.dirname(model_path), "lights.txt"))
    table = read_materials(os.path.splitext(model_path)[0] + ".mtl")
    mats = np.zeros(len(meshes), MATERIAL)
    verts, tris, tri_mesh, names = [], [], [], []
    base = 0
    for mi, (name, v, t) in enumerate(meshes):
        if name in table:
            a = table[name]
            mats[mi] = ((a[0], a[1], a[2]), a[3], a[4], a[5], a[6])
        else:
            mats[mi] = ((1.0, 1.0, 1.0), 1.0, 0.0, 0.0, 0.0)
        verts.append(v)
        tris.append(t + base)
        tri_mesh.append(np.full(len(t), mi, np.int32))
        names.append(name)
        base += len(v)
    return dict(
        verts=np.concatenate(verts).astype(np.float32) if verts else np.zeros((0, 3), np.float32),
        tris=np.concatenate(tris).astype(np.int32) if tris else np.zeros((0, 3), np.int32),
        tri_mesh=np.concatenate(tri_mesh) if tri_mesh else np.zeros(0, np.int32),
        mats=mats,
        lights=lights,
        light_shapes=read_light_shapes(os.path.join(os.path.dirname(model_path), "lights.txt")),
        names=names,
    )

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

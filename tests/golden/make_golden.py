"""Generates the committed golden vectors from the REFERENCE's own code (oracle/_ref, compiled from
/root/reference by `make -C oracle ref`). Run in the build container:  python tests/golden/make_golden.py

CPU part (runs here):   kdtree_ref.npz  — cukd::buildTree_host output for fixed inputs
                        lcg_kat.json    — owl::LCG<4> draws, owl normalize
                        view_ref.npz    — the photon viewer's glm projection (photon-viewer/src/hostCode.cu:53-75)
GPU part (--gpu, run on the B200 box through gpurun; writes into gpurun_out/ to travel back):
                        knn_ref.npz     — cukd::stackBased::knn<HeapCandidateList<k>> results
"""
import ctypes
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import PHOTON44, dup_cloud, unique_cloud, vp  # noqa: E402

ref = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libcukd_ref.so"))
ref.ref_lcg.restype = ctypes.c_uint32


def tree_case(n, seed):
    a = unique_cloud(n, seed)
    a["dir"][:, 0] = np.arange(n, dtype=np.float32)  # input index rides along
    t = a.copy()
    b = np.zeros(6, np.float32)
    ref.ref_build_tree_host(vp(t), n, vp(b))
    return a["pos"].copy(), t["dir"][:, 0].astype(np.int32), t["split_dim"].copy(), b


def cpu():
    out = {}
    for n, seed in [(1, 1), (2, 2), (3, 3), (10, 10), (1000, 1000), (4097, 4097)]:
        pos, order, dims, b = tree_case(n, seed)
        out[f"pos_{n}"] = pos
        out[f"order_{n}"] = order
        out[f"dims_{n}"] = dims
        out[f"bounds_{n}"] = b
    np.savez_compressed(os.path.join(HERE, "kdtree_ref.npz"), **out)
    kat = {"lcg": [], "normalize": []}
    for v0, v1 in [(0, 0), (7, 0), (1, 0), (123456, 0), (511, 255), (0xFFFFFFFF, 3)]:
        buf = np.zeros(8, np.float32)
        st = ref.ref_lcg(ctypes.c_uint32(v0), ctypes.c_uint32(v1), 8, vp(buf))
        kat["lcg"].append({"v0": v0, "v1": v1, "draws_bits": [int(x) for x in buf.view(np.uint32)], "state": int(st)})
    for v in [(1, 2, 3), (0.5, -0.25, 8.0), (1e-3, 1e-3, 1e-3)]:
        buf = np.zeros(3, np.float32)
        ref.ref_normalize(ctypes.c_float(v[0]), ctypes.c_float(v[1]), ctypes.c_float(v[2]), vp(buf))
        kat["normalize"].append({"v": list(v), "out_bits": [int(x) for x in buf.view(np.uint32)]})
    with open(os.path.join(HERE, "lcg_kat.json"), "w") as f:
        json.dump(kat, f, indent=1)
    print("wrote kdtree_ref.npz, lcg_kat.json")
    view()


VIEW_CASES = [((80, 30, 0), (10, 20, 0), (0, 1, 0), 0.87, 512, 512),          # config.toml.example camera
              ((1.5, -2, 7), (0.3, 0.2, 0.1), (0.1, 1, 0.05), 1.2, 1920, 1080),
              ((-30, 55, 20), (4, 4, 4), (0, 0, 1), 0.5, 333, 777)]


def view():
    """glm::lookAt / glm::perspective / mat4 * vec4 and the pixel rule, run with the reference's own glm."""
    g = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libglm_ref.so"))
    rng = np.random.default_rng(2024)
    out = {}
    for i, (frm, at, up, fovy, W, H) in enumerate(VIEW_CASES):
        frm, at, up = (np.array(v, np.float32) for v in (frm, at, up))
        m = np.zeros(16, np.float32)
        g.ref_glm_matrix(vp(frm), vp(at), vp(up), ctypes.c_float(fovy), W, H, vp(m))
        n = 6000
        pos = (rng.random((n, 3), dtype=np.float32) * np.float32(120) - np.float32(40)).astype(np.float32)
        pix = np.zeros((n, 2), np.int32)
        g.ref_glm_project(vp(frm), vp(at), vp(up), ctypes.c_float(fovy), W, H, vp(pos), ctypes.c_int64(n), vp(pix))
        out[f"cam_{i}"] = np.concatenate([frm, at, up, np.array([fovy, W, H], np.float32)])
        out[f"matrix_{i}"] = m
        out[f"pos_{i}"] = pos
        out[f"pix_{i}"] = pix
    np.savez_compressed(os.path.join(HERE, "view_ref.npz"), **out)
    print("wrote view_ref.npz")


def gpu():
    import torch

    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    res = {}
    for n, seed, k, cutoff in [(5000, 5, 50, 100.0), (5000, 6, 64, 100.0), (300, 7, 50, 100.0), (40, 8, 50, 100.0),
                               (5000, 9, 50, 0.6), (2000, 10, 8, 100.0)]:
        a = unique_cloud(n, seed)
        ref.ref_build_tree_host(vp(a), n, None)
        rng = np.random.default_rng(seed + 100)
        ext = float(n) * 2.0**-8
        q = (rng.random((64, 3), dtype=np.float32) * np.float32(ext * 1.2) - np.float32(0.1 * ext)).astype(np.float32)
        d_nodes = torch.from_numpy(a.view(np.uint8)).cuda()
        d_q = torch.from_numpy(q).cuda()
        d_out = torch.zeros((64, k), dtype=torch.int64, device="cuda")
        d_r2 = torch.zeros(64, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        rc = ref.ref_knn_gpu(ctypes.c_void_p(d_nodes.data_ptr()), n, ctypes.c_void_p(d_q.data_ptr()), 64, k,
                             ctypes.c_float(cutoff), ctypes.c_void_p(d_out.data_ptr()), ctypes.c_void_p(d_r2.data_ptr()))
        assert rc == 0, rc
        torch.cuda.synchronize()
        tag = f"n{n}_s{seed}_k{k}"
        res[f"tree_{tag}"] = a.view(np.uint8).reshape(n, 44)
        res[f"queries_{tag}"] = q
        res[f"entries_{tag}"] = d_out.cpu().numpy().view(np.uint64)  # heap order, as the reference returns them
        res[f"r2_{tag}"] = d_r2.cpu().numpy()
        res[f"cutoff_{tag}"] = np.float32(cutoff)
    np.savez_compressed(os.path.join(out_dir, "knn_ref.npz"), **res)
    print("wrote gpurun_out/knn_ref.npz")


if __name__ == "__main__":
    gpu() if "--gpu" in sys.argv else cpu()

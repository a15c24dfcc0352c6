"""CPU suite (-m "not gpu"): the oracle against the reference's golden vectors, the host logic, and
that the C-ABI library loads and exports every symbol include/pmhot.h declares."""
import ctypes
import json
import os
import re

import numpy as np
import pytest

from helpers import GOLDEN, PHOTON44, ROOT, dup_cloud, entries, oracle_brute, oracle_build, oracle_knn, unique_cloud, vp


# ---- ABI -------------------------------------------------------------------------------------
def test_abi_exports_every_declared_symbol(pkg):
    hdr = open(os.path.join(ROOT, "include", "pmhot.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(pm_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(pkg.ABI_SYMBOLS), declared ^ set(pkg.ABI_SYMBOLS)
    L = pkg.lib()  # raises if the .so is missing or lacks a symbol
    for name in declared:
        assert hasattr(L, name)
    assert L.pm_version() >= 100


def test_record_layouts(pkg):
    assert pkg.PHOTON40.itemsize == 40 and pkg.PHOTON44.itemsize == 44
    assert pkg.PHOTON40.fields["power"][1] == 24 and pkg.PHOTON40.fields["color"][1] == 28  # photon-mapping/include/photon.h
    assert pkg.PHOTON44.fields["power"][1] == 36 and pkg.PHOTON44.fields["split_dim"][1] == 43  # ray-tracer/include/photon.h
    assert pkg.MATERIAL.itemsize == 28


def test_no_cpu_fallback_without_gpu(pkg):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.PmError) as e:
        pkg.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_touch_oracle():
    """Nothing under photon-mapping_b200/ may import, link or execute oracle/."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "photon-mapping_b200")):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh", "Makefile")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "libpm_oracle" not in txt and "oracle/" not in txt and "orc_" not in txt, os.path.join(dirpath, f)


# ---- golden vectors from the reference's own code ---------------------------------------------
def test_lcg_against_reference_kat(orc):
    kat = json.load(open(os.path.join(GOLDEN, "lcg_kat.json")))
    orc.orc_lcg_draws.restype = ctypes.c_uint32
    for case in kat["lcg"]:
        buf = np.zeros(8, np.float32)
        st = orc.orc_lcg_draws(ctypes.c_uint32(case["v0"]), ctypes.c_uint32(case["v1"]), 8, vp(buf))
        assert [int(x) for x in buf.view(np.uint32)] == case["draws_bits"]
        assert st == case["state"]
    for case in kat["normalize"]:
        buf = np.zeros(3, np.float32)
        v = case["v"]
        orc.orc_normalize(ctypes.c_float(v[0]), ctypes.c_float(v[1]), ctypes.c_float(v[2]), vp(buf))
        assert [int(x) for x in buf.view(np.uint32)] == case["out_bits"]


def test_survey_kats(orc):
    """SURVEY §8c self-derived KATs: LCG<4>.init(0,0) → 0.294449925, 0.695515215, state 4105309513."""
    orc.orc_lcg_draws.restype = ctypes.c_uint32
    buf = np.zeros(2, np.float32)
    st = orc.orc_lcg_draws(0, 0, 2, vp(buf))
    assert buf[0] == np.float32(0.294449925) and buf[1] == np.float32(0.695515215) and st == 4105309513
    buf1 = np.zeros(1, np.float32)
    orc.orc_lcg_draws(7, 0, 1, vp(buf1))
    assert buf1[0] == np.float32(0.883485258)


@pytest.mark.parametrize("n", [1, 2, 3, 10, 1000, 4097])
def test_tree_against_reference_golden(orc, n):
    g = np.load(os.path.join(GOLDEN, "kdtree_ref.npz"))
    a = np.zeros(n, PHOTON44)
    a["pos"] = g[f"pos_{n}"]
    a["dir"][:, 0] = np.arange(n, dtype=np.float32)
    a["split_dim"] = 255
    t, b = oracle_build(orc, a)
    assert np.array_equal(t["dir"][:, 0].astype(np.int32), g[f"order_{n}"])
    assert np.array_equal(t["split_dim"], g[f"dims_{n}"])
    assert np.array_equal(b, g[f"bounds_{n}"])


@pytest.mark.parametrize("n", [1, 2, 5, 31, 32, 33, 1000, 20000])
def test_tree_against_reference_live(orc, ref, n):
    """byte-identical arrays vs cukd::buildTree_host (the check testBuildersSameResult.cu makes)."""
    a = unique_cloud(n, n + 17)
    t, b = oracle_build(orc, a)
    r = a.copy()
    rb = np.zeros(6, np.float32)
    ref.ref_build_tree_host(vp(r), n, vp(rb))
    assert t.tobytes() == r.tobytes()
    assert np.array_equal(b, rb)


def test_knn_against_reference_golden(orc):
    path = os.path.join(GOLDEN, "knn_ref.npz")
    if not os.path.exists(path):
        pytest.skip("knn_ref.npz not generated yet (make_golden.py --gpu on the B200 box)")
    g = np.load(path)
    tags = [k[len("tree_"):] for k in g.files if k.startswith("tree_")]
    assert tags
    for tag in tags:
        tree = np.ascontiguousarray(g[f"tree_{tag}"]).view(PHOTON44).reshape(-1)
        q = g[f"queries_{tag}"]
        want = g[f"entries_{tag}"]
        k = want.shape[1]
        cutoff = float(g[f"cutoff_{tag}"])
        for i in range(len(q)):
            heap = np.empty(k, np.uint64)
            r2 = orc.orc_knn_traverse(vp(tree), len(tree), vp(np.ascontiguousarray(q[i])), k, cutoff, vp(heap), None)
            assert np.array_equal(heap, want[i]), (tag, i)  # same entries in the same heap order
            assert np.float32(r2) == g[f"r2_{tag}"][i]


# ---- oracle self-consistency -------------------------------------------------------------------
def test_subtree_sizes(orc):
    for n in [1, 2, 3, 6, 7, 8, 100, 1023, 1024, 1025]:
        def rec(i):
            return 0 if i >= n else 1 + rec(2 * i + 1) + rec(2 * i + 2)
        for root in range(min(n, 40)):
            assert orc.orc_subtree_size(root, n) == rec(root)


@pytest.mark.parametrize("n,k", [(1, 8), (7, 8), (49, 50), (50, 50), (51, 50), (3000, 50), (3000, 64)])
def test_traversal_equals_bruteforce(orc, n, k):
    a = unique_cloud(n, n * 3 + k)
    tree, _ = oracle_build(orc, a)
    assert orc.orc_kd_check(vp(tree), n) == 1
    rng = np.random.default_rng(5)
    ext = n * 2.0**-8
    q = (rng.random((50, 3), dtype=np.float32) * np.float32(ext)).astype(np.float32)
    for cutoff in (100.0, max(ext * 0.1, 0.01)):
        ids, d2, r2 = oracle_knn(orc, tree, q, k, cutoff)
        e = entries(ids, d2)
        for i in range(len(q)):
            want, wr2, tie = oracle_brute(orc, tree, q[i], k, cutoff)
            if tie:
                continue
            assert np.array_equal(e[i], want)
            assert r2[i] == np.float32(wr2)


def test_tree_with_duplicates_is_valid_and_index_tiebroken(orc):
    n = 5000
    a = dup_cloud(n, 3)
    t, _ = oracle_build(orc, a)
    assert orc.orc_kd_check(vp(t), n) == 1
    assert sorted(t["dir"][:, 0].astype(int).tolist()) == list(range(n))


def test_gather_matches_python_restatement(orc):
    """gatherPhotons (ray-tracer/cuda/shading.h:93-121) restated in numpy float32."""
    n, k = 400, 50
    a = unique_cloud(n, 77)
    tree, _ = oracle_build(orc, a)
    q = np.array([[0.7, 0.4, 0.9]], np.float32)
    brdf = np.array([0.25], np.float32)
    out = np.zeros((1, 3), np.float32)
    orc.orc_gather_batch(vp(tree), n, vp(q), vp(brdf), 1, k, 100.0, vp(out), 1)
    heap = np.empty(k, np.uint64)
    r2 = np.float32(orc.orc_knn_traverse(vp(tree), n, vp(q[0]), k, 100.0, vp(heap), None))
    f32 = np.float32
    flux = np.zeros(3, f32)
    for e in heap:
        pid = int(e & np.uint64(0xFFFFFFFF))
        if pid >= 2**31:
            continue
        p = tree[pid]
        d = p["pos"] - q[0]
        dist = np.sqrt(f32(f32(d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]))
        w = f32(1) - f32(f32(dist / np.sqrt(r2)) * f32(1.1))
        flux = (flux + f32(f32(brdf[0] * p["power"]) * w) * p["color"]).astype(f32)
    denom = f32(f32(f32(f32(1) - f32(f32(2.0) / f32(3.0)) * f32(f32(1.0) / f32(1.1))) * f32(2)) * f32(3.141592653)) * r2
    assert np.array_equal(out[0], (flux / denom).astype(f32))

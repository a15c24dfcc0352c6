"""CPU tests of the C++ hosts (photon-mapping_b200/host): config.toml reader, glb / lights / mtl import
(≙ common/src/configLoader.h, common/src/assetImporter.cxx) through `photonMapping --dump-scene`, which
needs no GPU, compared with the Python importer; PNG writer round trip is covered by the GPU host test."""
import os
import subprocess

import numpy as np
import pytest

from helpers import HOST_DIR, load_scene, write_config, write_scene_files


@pytest.fixture(scope="module")
def hosts():
    exe = os.path.join(HOST_DIR, "photonMapping")
    if not os.path.exists(exe) or not os.path.exists(os.path.join(HOST_DIR, "rayTracer")):
        env = {k: v for k, v in os.environ.items() if k not in ("CC", "CXX")}
        subprocess.run(["make"], cwd=HOST_DIR, check=True, env=env)
    return HOST_DIR


def _dump(hosts, cfg):
    r = subprocess.run([os.path.join(hosts, "photonMapping"), cfg, "--dump-scene"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    meshes, lights, pm = [], [], None
    for ln in r.stdout.splitlines():
        t = ln.split()
        if not t:
            continue
        if t[0] == "mesh":
            name = ln.split('"')[1]
            rest = ln.split('"')[2].split()
            meshes.append(dict(name=name, v=[], f=[], mat=[float(x) for x in rest[3:10]]))
        elif t[0] == "v":
            meshes[-1]["v"].append([float(x) for x in t[1:]])
        elif t[0] == "f":
            meshes[-1]["f"].append([int(x) for x in t[1:]])
        elif t[0] == "light":
            lights.append([float(x) for x in t[1:]])
        elif t[0] == "photon-mapper":
            pm = [int(x) for x in t[1:]]
    return meshes, lights, pm


def test_scene_import_matches_python_importer(hosts, pkg, tmp_path):
    import importlib.util

    sc = load_scene("cornell_box")
    rot = [0.0, float(np.sin(0.3)), 0.0, float(np.cos(0.3))]
    tr = [None] * len(sc["mats"])
    tr[0] = {"translation": [1.0, -2.0, 0.5], "scale": [1.0, 2.0, 1.0]}
    tr[1] = {"rotation": rot}
    glb = write_scene_files(str(tmp_path), sc, "box", tr)
    cfg = str(tmp_path / "config.toml")
    write_config(cfg, glb, str(tmp_path / "g.txt"), str(tmp_path / "c.txt"), str(tmp_path / "o.png"))
    meshes, lights, pm = _dump(hosts, cfg)
    spec = importlib.util.spec_from_file_location("pm_scene_io", os.path.join(os.path.dirname(HOST_DIR), "scene_io.py"))
    sio = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sio)
    want = sio.import_scene(glb)
    assert pm == [20000, 10000, 6]  # 20_000: TOML digit separators
    assert len(meshes) == len(want["names"]) and [m["name"] for m in meshes] == want["names"]
    v = np.concatenate([np.array(m["v"], np.float32).reshape(-1, 3) for m in meshes])
    base, f = 0, []
    for m in meshes:
        f.append(np.array(m["f"], np.int32).reshape(-1, 3) + base)
        base += len(m["v"])
    assert np.array_equal(np.concatenate(f), want["tris"])
    assert v.shape == want["verts"].shape and np.abs(v - want["verts"]).max() <= 1e-5 * np.abs(want["verts"]).max()
    for m, w in zip(meshes, want["mats"]):
        assert np.allclose(m["mat"], [*w["albedo"], w["diffuse"], w["specular"], w["transmission"], w["refraction_idx"]], rtol=1e-6)
    assert np.allclose(np.array(lights), [[*l["pos"], *l["rgb"], l["power"]] for l in want["lights"]])


def test_missing_material_file_warns_and_missing_lights_fails(hosts, tmp_path):
    sc = load_scene("cornell_box")
    glb = write_scene_files(str(tmp_path), sc, "box")
    cfg = str(tmp_path / "config.toml")
    write_config(cfg, glb, "g.txt", "c.txt", "o.png")
    os.remove(tmp_path / "box.mtl")
    meshes, _, _ = _dump(hosts, cfg)  # assetImporter.cxx:149-152: only a warning, white diffuse default
    assert all(m["mat"] == [1, 1, 1, 1, 0, 0, 0] for m in meshes)
    os.remove(tmp_path / "lights.txt")
    r = subprocess.run([os.path.join(hosts, "photonMapping"), cfg, "--dump-scene"], capture_output=True, text=True)
    assert r.returncode != 0 and "Unable to open file" in r.stderr  # assetImporter.cxx:108-110


def test_config_errors(hosts, tmp_path):
    cfg = tmp_path / "config.toml"
    cfg.write_text("[data]\nphotons_file = \"a\"\n")
    r = subprocess.run([os.path.join(hosts, "photonMapping"), str(cfg), "--dump-scene"], capture_output=True, text=True)
    assert r.returncode != 0 and "missing" in r.stderr
    r = subprocess.run([os.path.join(hosts, "rayTracer"), str(tmp_path / "nope.toml")], capture_output=True, text=True)
    assert r.returncode != 0 and "cannot open config" in r.stderr


def test_photon_file_formats_roundtrip(pkg, tmp_path):
    """text (≙ writeAlivePhotons / readPhotonsFromFile: 9 floats, 6 decimals) and the lossless binary hand-off"""
    rng = np.random.default_rng(5)
    p = np.zeros(1000, pkg.PHOTON40)
    p["pos"] = (rng.random((1000, 3)) * 80 - 40).astype(np.float32)
    p["dir"] = rng.standard_normal((1000, 3)).astype(np.float32)
    p["color"] = rng.random((1000, 3)).astype(np.float32)
    txt, binf = str(tmp_path / "p.txt"), str(tmp_path / "p.bin")
    pkg.write_photons_text(txt, p)
    first = open(txt).readline().split()
    assert len(first) == 9 and all(len(t.split(".")[1]) == 6 for t in first)  # std::fixed, setprecision(6)
    back = pkg.read_photons_text(txt)
    assert len(back) == 1000 and np.abs(back["pos"] - p["pos"]).max() <= 1e-5 and (back["power"] == 0).all()
    pkg.write_photons_bin(binf, p)
    assert os.path.getsize(binf) == 16 + 40 * 1000
    assert pkg.read_photons_bin(binf).tobytes() == p.tobytes()
    pkg.write_photons_bin(binf, p[:0])
    assert len(pkg.read_photons_bin(binf)) == 0
    with pytest.raises(pkg.PmError):
        pkg.read_photons_bin(txt)  # wrong magic
    with pytest.raises(pkg.PmError):
        pkg.read_photons_text(str(tmp_path / "missing.txt"))

"""GPU test of the C++ hosts: `photonMapping` then `rayTracer` on a glb scene, exactly the reference's
two-executable workflow (config.toml → photon text files → PNG), compared with the same path driven
through the ctypes binding."""
import os
import subprocess

import numpy as np
import pytest

from helpers import HOST_DIR, load_scene, read_png_rgba8, write_config, write_scene_files

pytestmark = pytest.mark.gpu


def test_two_pass_workflow_matches_python_host(pkg, ctx, tmp_path):
    import importlib.util

    for exe in ("photonMapping", "rayTracer", "photonViewer"):
        assert os.path.exists(os.path.join(HOST_DIR, exe)), "hosts are built by __graft_entry__.build()"
    sc0 = load_scene("cornell_box")
    glb = write_scene_files(str(tmp_path), sc0, "box")
    W, H = 96, 64
    cfg = str(tmp_path / "config.toml")
    gfile, cfile, png = str(tmp_path / "global.txt"), str(tmp_path / "caustic.txt"), str(tmp_path / "result.png")
    write_config(cfg, glb, gfile, cfile, png, W=W, H=H, spp=1, depth=1, max_depth=6, casted=(20000, 10000))
    r = subprocess.run([os.path.join(HOST_DIR, "photonMapping"), cfg], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    r = subprocess.run([os.path.join(HOST_DIR, "rayTracer"), cfg], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    assert "Loaded" in r.stdout and "Photon maps built" in r.stdout

    # same path through the Python host, on the scene as the Python importer reads it
    spec = importlib.util.spec_from_file_location("pm_scene_io", os.path.join(os.path.dirname(HOST_DIR), "scene_io.py"))
    sio = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sio)
    sc = sio.import_scene(glb)
    scene = pkg.Scene(ctx, sc["verts"], sc["tris"], sc["tri_mesh"], sc["mats"])
    g40, ng, _, _ = scene.emit_photons(sc["lights"], 20000, 6, False)
    c40, nc, _, _ = scene.emit_photons(sc["lights"], 10000, 6, True)
    assert ng > 1000
    got_g, got_c = pkg.read_photons_text(gfile), pkg.read_photons_text(cfile)
    assert len(got_g) == ng and len(got_c) == nc
    pkg.write_photons_text(str(tmp_path / "g2.txt"), g40)
    assert open(gfile).read() == open(tmp_path / "g2.txt").read()  # identical text files
    # render from the text files (6-decimal photons), as rayTracer does
    g44, c44 = pkg.photons_to_maps(ctx, got_g, got_c)
    gmap = pkg.PhotonMap.build(ctx, g44)
    cmap = pkg.PhotonMap.build(ctx, c44) if len(c44) else None
    cam = pkg.make_camera((80.0, 30.0, 0.0), (10.0, 20.0, 0.0), (0.0, 1.0, 0.0), 0.87, W, H)
    fb = scene.render(sc["lights"], gmap, cmap, cam, W, H, 1, 1)
    img = read_png_rgba8(png)
    assert img.shape == (H, W)
    # the C++ host computes the camera with std::cos, the Python host with numpy: allow an 8-bit step
    a, b = img.view(np.uint8).reshape(H, W, 4).astype(int), fb.view(np.uint8).reshape(H, W, 4).astype(int)
    assert (np.abs(a - b) <= 1).mean() > 0.995 and (a[..., 3] == b[..., 3]).all()
    assert (img[1:] != 0).any()
    # third executable: the photon viewer on the same files (≙ photonViewer), against the binding
    r = subprocess.run([os.path.join(HOST_DIR, "photonViewer"), cfg], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    for path, photons in ((png + ".viewer.png", got_g), (png + ".caustics-viewer.png", got_c)):
        want = scene.view_photons(photons, (80.0, 30.0, 0.0), (10.0, 20.0, 0.0), (0.0, 1.0, 0.0), 0.87, W, H)
        assert np.array_equal(read_png_rgba8(path), want)
    assert (read_png_rgba8(png + ".viewer.png") != 0xFF000000).sum() > 500


def test_two_rank_hosts_write_the_single_process_files(tmp_path):
    """Multi-process mode of the C++ hosts (pm_ranks.hpp): two ranks on two GPUs — sharded emission, the sort-free
    NCCL exchange, the sharded map build, pixel ranges merged by one all-reduce — produce byte-identical photon files
    and the identical PNG of the single-process run. Needs two GPUs (skipped on the driver's one-GPU test box;
    run with `gpurun --gpus 2`, record in profiles/)."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    sc0 = load_scene("cornell_box")
    glb = write_scene_files(str(tmp_path), sc0, "box")
    W, H = 160, 96
    out = {}
    for tag, world in (("one", 1), ("two", 2)):
        d = tmp_path / tag
        d.mkdir()
        cfg = str(d / "config.toml")
        gfile, cfile, png = str(d / "global.bin"), str(d / "caustic.txt"), str(d / "result.png")
        write_config(cfg, glb, gfile, cfile, png, W=W, H=H, spp=1, depth=1, max_depth=6, casted=(60001, 20000))
        for exe in ("photonMapping", "rayTracer"):
            procs = []
            for r in range(world):
                cmd = [os.path.join(HOST_DIR, exe), cfg]
                if world > 1:
                    cmd += ["--rank", str(r), "--world", str(world), "--comm-file", str(d / f"{exe}.id"), "--device", str(r)]
                procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
            for p in procs:
                o, _ = p.communicate(timeout=300)
                assert p.returncode == 0, o
        out[tag] = (open(gfile, "rb").read(), open(cfile, "rb").read(), read_png_rgba8(png))
    assert len(out["one"][0]) > 1_000_000
    assert out["one"][0] == out["two"][0]  # binary global photon file
    assert out["one"][1] == out["two"][1]  # text caustic photon file
    assert np.array_equal(out["one"][2], out["two"][2]) and (out["one"][2][1:] != 0).any()

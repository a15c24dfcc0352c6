"""Extracts the reference's Cornell-box, sphere and dragon-box assets into small committed fixtures (geometry, materials,
lights), so the tests and the bench can use the reference's own C1 scene on the GPU box where
/root/reference does not exist.  Run in the build container:  python tests/golden/make_scenes.py
Source: /root/reference/assets/models/cornell-box/{cornell-box.glb, cornell-box.mtl, lights.txt}."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

pkg = graft.load_package()
from photon_mapping_b200 import scene_io  # noqa: E402

for name, rel in [("cornell_box", "cornell-box/cornell-box.glb"), ("sphere", "sphere/sphere.glb"),
                  ("dragon_box", "dragon/dragon-box.glb")]:
    path = os.path.join("/root/reference/assets/models", rel)
    if not os.path.exists(path):
        print("skip", path)
        continue
    s = scene_io.import_scene(path)
    print(name, "verts", s["verts"].shape, "tris", s["tris"].shape, "meshes", len(s["names"]), s["names"], "lights", len(s["lights"]))
    np.savez_compressed(os.path.join(HERE, f"scene_{name}.npz"), verts=s["verts"], tris=s["tris"], tri_mesh=s["tri_mesh"],
                        mats=s["mats"].view(np.uint8), lights=s["lights"].view(np.uint8), names=np.array(s["names"]))

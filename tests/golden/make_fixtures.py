"""Generates the committed fixtures under tests/golden/ from the reference tree (run in the build container only;
/root/reference does not exist on the GPU box).

  cbox_scene.npz  -- resources/data/scenes/cbox/cbox-rgb.xml flattened by mitsuba2_b200.xml (meshes with to_world
                     baked + computed vertex normals, BSDFs, area emitter, sensor), max_depth 6.
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from mitsuba2_b200 import xml as X   # noqa: E402

REF = os.environ.get("MTS_REFERENCE", "/root/reference")

if __name__ == "__main__":
    sd = X.load_file(os.path.join(REF, "resources/data/scenes/cbox/cbox-rgb.xml"), spp=64, res=256)
    for m in sd.meshes:
        m.name = os.path.basename(m.name)
    sd.save_npz(os.path.join(HERE, "cbox_scene.npz"))
    print("cbox:", len(sd.meshes), "meshes,", sum(m.faces.shape[0] for m in sd.meshes), "triangles")

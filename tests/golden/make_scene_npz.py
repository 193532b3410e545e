"""Writes the flat records of the built-in Cornell box (BASELINE.json configs[0]/[1]) as tests/golden/scene_cornell.npz, so
that `bench.py --impl reference` can feed the CPU oracle without importing the product package, and those of the reference's
own staircase2/scene.xml (read by the host layer's Mitsuba loader from /root/reference: OBJ shapes, roughconductor,
roughplastic, dielectric, twosided) as tests/golden/scene_staircase2.npz, so that the GPU parity test of that scene can
run on the GPU box, where the reference checkout does not exist.

    python tests/golden/make_scene_npz.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import gpuspectral_b200 as g          # host layer only: scene construction
from oracle import oracle as orc

if __name__ == "__main__":
    out = os.path.join(ROOT, "tests", "golden", "scene_cornell.npz")
    orc.OracleScene.dump_host_scene(g.Scene.builtin("cornell"), out)
    print(out, os.path.getsize(out), "bytes")
    xml = "/root/reference/src/GPUSpectral/assets/scenes/staircase2/scene.xml"
    if os.path.exists(xml):
        sc = g.Scene.load_xml(xml)
        out = os.path.join(ROOT, "tests", "golden", "scene_staircase2.npz")
        orc.OracleScene.dump_host_scene(sc, out, max_texture=256)     # its three JPEG textures box-filtered to <= 256 pixels
        print(out, os.path.getsize(out), "bytes;", sc.instanced_triangles, "triangles; loader warnings:", sc.warnings)

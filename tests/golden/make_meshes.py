"""Extract the small 3D macro meshes of the reference (data/meshes/3D/*.msh) into meshes.json.

Run in the build container (needs /root/reference); the JSON travels to the GPU box.
"""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from hyteg_b200.meshes import from_gmsh  # noqa: E402

REF = "/root/reference/data/meshes/3D"
NAMES = ["cube_6el", "cube_24el", "cube_6el_offcenter", "pyramid_4el", "pyramid_4el_offcenter", "pyramid_tilted_4el",
         "regular_octahedron_8el", "tet_tilted_1el", "three_tets_with_two_joint_faces", "cube_center_at_origin_24el"]
out = {}
for n in NAMES:
    v, c = from_gmsh(os.path.join(REF, n + ".msh"))
    out[n] = {"vertices": v.tolist(), "cells": c.tolist()}
with open(os.path.join(os.path.dirname(__file__), "meshes.json"), "w") as f:
    json.dump(out, f)
print({k: (len(v["vertices"]), len(v["cells"])) for k, v in out.items()})

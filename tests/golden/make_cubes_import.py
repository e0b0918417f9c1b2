#!/usr/bin/env python
"""Generates tests/golden/cubes_import_golden.npz: what the library's loader reads out of the reference's OWN serialised
worlds (apps/cubes/assets/worlds/main.object and shadow_test.object, with their .obj meshes), so that the device tests
can rebuild those worlds where /root/reference does not exist (the GPU box). Run here, where the reference tree is:
    python tests/golden/make_cubes_import.py
The arrays are checked against the restated loader and the SURVEY's transcription by tests/test_scene_import.py."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from orion_b200 import lib  # noqa: E402

REF_ASSETS = "/root/reference/apps/cubes/assets"
KEYS = ("parent", "local_trs", "local_aabb", "flags", "entity_node", "lights", "shading", "cameras")

out = {}
for name in ("main", "shadow_test"):
    s = lib.load_scene(os.path.join(REF_ASSETS, "worlds", name + ".object"), REF_ASSETS, 1440, 900)
    for k in KEYS:
        out[f"{name}_{k}"] = np.asarray(s[k])
    out[f"{name}_shadow_map_resolution"] = np.int32(s["shadow_map_resolution"])
np.savez_compressed(os.path.join(HERE, "cubes_import_golden.npz"), **out)
print({k: v.shape for k, v in out.items()})

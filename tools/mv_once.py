#!/usr/bin/env python
"""Runs configs[3]'s multi-view culls (camera + 4 cascades; 64 point lights x 6 faces) a few times, for ncu:
    ncu --set full -k regex:mv_ ... python tools/mv_once.py [entities]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import lib, scenes  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
s = scenes.config4(n=n)
w = lib.World(0)
w.reserve(n)
f = w.add_entities(s["aabb"], s["flags"])
w.set_world_aabbs(f, s["aabb"])
w.add_lights(s["lights"])
cam_views = np.concatenate([np.asarray(s["view"]).reshape(1), s["cascades"]])
cam_planes = np.stack([lib.view_planes(v).reshape(24) for v in cam_views])
light_ids = np.arange(len(s["lights"]), dtype=np.uint32)
for _ in range(3):
    off, _, _ = w.cull_views_device(cam_planes)
    ll, lv, loff, _, _ = w.cull_shadow_views_device(light_ids)
torch.cuda.synchronize()
print("camera lists", [int(off[i + 1] - off[i]) for i in range(len(cam_views))], "light face pairs", int(loff[-1]))
w.close()

#!/usr/bin/env python
"""Runs the single-view cull of the 12.5 M-entity shard a few times (for ncu: `OVIS_CULL_VARIANT=k ncu ... python tools/cull_once.py`)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import lib, scenes  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 12_500_000
s = scenes.config5(n=n, seed=4)
w = lib.World(0)
f = w.add_entities(s["aabb"], s["flags"])
w.set_world_aabbs(f, s["aabb"])
out = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
planes = lib.view_planes(s["view"])
for _ in range(4):
    w.cull_planes_device(planes, out.data_ptr() + 4, out.data_ptr())
torch.cuda.synchronize()
print("visible", int(out[0].item()))
w.close()

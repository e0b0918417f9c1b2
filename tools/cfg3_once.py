#!/usr/bin/env python
"""configs[2] (10 M entities under a transform tree, 4096 point lights): the frame's culls a few times, for ncu launch lists:
    ncu --metrics gpu__time_duration.sum -k regex:"mv_|cull_|update_" ... python tools/cfg3_once.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import lib, scenes  # noqa: E402

s = scenes.config3()
n_ent, n_nodes = s["n_entities"], s["n_nodes"]
w = lib.World(0)
w.reserve(n_ent, n_nodes)
e0 = w.add_entities(s["local_aabb"], s["flags"])
w.add_nodes(s["parent"], s["local_trs"])
w.bind_entities(s["entity_node"], np.arange(e0, e0 + n_ent, dtype=np.uint32))
w.add_lights(s["lights"])
w.update_transforms()
g0, g1 = s["group_range"]
group_ids = np.arange(g0, g1, dtype=np.uint32)
for f in range(1, 4):                                    # three frames of moved groups: the dirty-subtree update
    w.set_node_local_trs(group_ids, scenes.move_groups(s, f))
    w.update_transforms()
ents, ls = w.cull(s["view"])
shadow_ids = ls[(s["lights"]["flags"][ls] & 1) == 1]
for _ in range(3):
    ll, lv, off, _, _ = w.cull_shadow_views_device(shadow_ids)
torch.cuda.synchronize()
print("visible", len(ents), "lights", len(ls), "shadow views", len(ll), "pairs", int(off[-1]))
w.close()

#!/usr/bin/env python
"""configs[2]'s transform update (10 M nodes, every group moved) timed with CUDA events on the world's stream, under two
settings of an environment switch (A/B of a kernel variant on one box): python tools/ab_update.py [ENV_NAME]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orion_b200 import lib, scenes  # noqa: E402

s = scenes.config3()
n_ent, n_nodes = s["n_entities"], s["n_nodes"]
g0, g1 = s["group_range"]
group_ids = np.arange(g0, g1, dtype=np.uint32)
moves = [scenes.move_groups(s, f) for f in range(1, 5)]
out = {}
SWITCH = sys.argv[1] if len(sys.argv) > 1 else "OVIS_AB_SWITCH"
for mode in ("1", "0", "1", "0"):
    os.environ[SWITCH] = mode
    w = lib.World(0)
    stream = torch.cuda.Stream()
    w.set_stream(stream.cuda_stream)                       # the events below must sit on the stream the kernels run on
    w.reserve(n_ent, n_nodes)
    e0 = w.add_entities(s["local_aabb"], s["flags"])
    w.add_nodes(s["parent"], s["local_trs"])
    w.bind_entities(s["entity_node"], np.arange(e0, e0 + n_ent, dtype=np.uint32))
    w.update_transforms()
    torch.cuda.synchronize()
    whole, kernels = [], []
    for f in range(24):
        a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        a.record(stream)
        w.set_node_local_trs(group_ids, moves[f % 4])
        b.record(stream)
        w.update_transforms()
        c.record(stream)
        torch.cuda.synchronize()
        if f >= 4:
            whole.append(a.elapsed_time(c))
            kernels.append(b.elapsed_time(c))
    out.setdefault(mode, []).append((float(np.median(whole)), float(np.median(kernels)), float(np.min(kernels))))
    chk = w.world_aabbs(0, 1000).copy()
    out.setdefault("chk" + mode, []).append(float(np.nansum(chk)))
    w.close()
print(out)

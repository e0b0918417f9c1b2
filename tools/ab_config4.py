#!/usr/bin/env python
"""configs[3] (10 M entities, camera + 4 depth slices + 64 point lights x 6) with and without the shared-side-plane runs of
the dense pass (OVIS_MV_SHARED_SIDES): python tools/ab_config4.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench_configs  # noqa: E402

out = {}
for mode in ("1", "0"):
    os.environ["OVIS_MV_SHARED_SIDES"] = mode
    r = bench_configs.config4(False, full_parity=(mode == "1"), brute=False)
    out[mode] = {k: v for k, v in r.items() if k.endswith("_ms") or "parity" in k or "exact" in k}
print(json.dumps(out))

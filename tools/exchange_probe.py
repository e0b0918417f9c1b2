#!/usr/bin/env python
"""Where does the fused exchange lose time against the plain cull? One GPU, 12.5 M entities, back-to-back frames:

  plain      ovis_cull_planes_device
  dst1       exchange entry point, one local destination, no flags
  dst1+flags the same, waiting on / signalling its own flag (frame f waits for f-2): fence + done counter + flag store
  dstN       N local destinations (stores multiplied, no NVLink), with flags

    python tools/exchange_probe.py [entities]
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from orion_b200 import lib, scenes  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 12_500_000
    s = scenes.config5(n=n, seed=4)
    stream = torch.cuda.Stream()
    w = lib.World(0)
    w.set_stream(stream.cuda_stream)
    w.reserve(n, 0)
    first = w.add_entities(s["aabb"], s["flags"])
    w.set_world_aabbs(first, s["aabb"])
    planes = np.ascontiguousarray(lib.view_planes(s["view"]), dtype=np.float32).reshape(24)
    cap = n // 8
    with torch.cuda.stream(stream):
        bufs = [torch.zeros(cap + 1, dtype=torch.int32, device="cuda") for _ in range(8)]
        flags = torch.zeros(8, dtype=torch.int32, device="cuda")
        full = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
    stream.synchronize()
    out = {"entities": n}
    K = 200

    def run(name, fn):
        with torch.cuda.stream(stream):
            for f in range(1, 6):
                fn(f)
            stream.synchronize()
            flags.zero_()
            stream.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for f in range(1, K + 1):
                fn(f)
            e1.record(stream)
        stream.synchronize()
        out[name] = e0.elapsed_time(e1) / K * 1e3

    run("plain_us", lambda f: w.cull_planes_device(planes, full.data_ptr() + 4, full.data_ptr()))
    for nd in (1, 2, 4, 8):
        outs = [b.data_ptr() + 4 for b in bufs[:nd]]
        cnts = [b.data_ptr() for b in bufs[:nd]]
        plan0 = w.exchange_plan(outs, cnts, [])
        run(f"dst{nd}_noflags_us", lambda f: w.cull_exchange_planned(planes, plan0, cap, 0, 0, 0, 0))
        sig = [flags.data_ptr() + 4 * k for k in range(nd)]
        plan1 = w.exchange_plan(outs, cnts, sig)
        run(f"dst{nd}_flags_us", lambda f: w.cull_exchange_planned(planes, plan1, cap, flags.data_ptr(), nd, max(f - 2, 0), f))
    out["exchange_error"] = w.exchange_error()
    w.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Per-call cost of vmn_intersect_static (device scratch in the library's static array, no allocation) against vmn_intersect (the
context's grow-only cudaMalloc buffers) on the C2 scene, for the batch sizes a serving loop sends: host rays in, host hits out, wall
clock per call (median of `reps`), plus the kernel's own device time from vmn_get_stats.
  python tools/bench_static.py [reps] > profiles/<round>_static_query.json"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]


def main():
    import common
    from vermilion_b200 import engine, scenes
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    pos, nrm, uv = scenes.spheres(4, 3, 6, 150.0)  # C2: 983 040 triangles
    box = (np.array([-1900, 10, -1900], np.float32), np.array([1900, 990, 1900], np.float32))
    rays_all = common.aimed_rays(pos, engine.STATIC_MAX_RAYS, 7, spread=0.05, origin_box=box)
    out = {"scene": "C2, 983040 triangles", "reps": reps, "rows": []}
    with engine.Engine(device=0) as e:
        e.set_scene(pos, nrm, uv)
        for n in (1, 256, 4096, 65536):
            rays = np.ascontiguousarray(rays_all[:n])
            row = {"rays": n}
            ref = None
            for name, fn in (("vmn_intersect", e.intersect), ("vmn_intersect_static", e.intersect_static)):
                for _ in range(5):
                    h = fn(rays)
                t, k = [], []
                for _ in range(reps):
                    t0 = time.perf_counter()
                    h = fn(rays)
                    t.append(time.perf_counter() - t0)
                    k.append(e.stats()["extend_ms"])
                if ref is None:
                    ref = h
                row[name] = {"call_us_median": 1e6 * float(np.median(t)), "call_us_p99": 1e6 * float(np.percentile(t, 99)),
                             "kernel_us_median": 1e3 * float(np.median(k)), "mrays_per_s": n / float(np.median(t)) / 1e6}
                row["identical"] = bool(np.array_equal(h.view(np.uint8), ref.view(np.uint8)))
            out["rows"].append(row)
    print(json.dumps(out))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Per-scene traversal statistics on one GPU: nodes / triangles visited per ray (counting build), Mrays/s, per-stage times.
  python tools/scene_stats.py city:24 spheres soup:1000000 [--res 1920x1080] [--spp 8]"""
import argparse, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vermilion_b200 import engine, scenes

ap = argparse.ArgumentParser()
ap.add_argument("scenes", nargs="+")
ap.add_argument("--res", default="1920x1080")
ap.add_argument("--spp", type=int, default=8)
args = ap.parse_args()
w, h = (int(x) for x in args.res.split("x"))
for name in args.scenes:
    kind, _, arg = name.partition(":")
    t0 = time.time()
    if kind == "city":
        pos, nrm, uv = scenes.city(64, 64, int(arg or 24), 2024); cam = ((0.0, 900.0, 1950.0), (25.0, 0.0, 0.0))
    elif kind == "spheres":
        pos, nrm, uv = scenes.spheres(4, 3, 6, 150.0); cam = ((0.0, 500.0, 1900.0), (0.0, 0.0, 0.0))
    elif kind == "cornell":
        pos, nrm, uv = scenes.cornell(); cam = ((0.0, 275.0, 1100.0), (0.0, 0.0, 0.0))
    else:
        raise SystemExit(f"unknown scene {name}")
    out = {"scene": name, "triangles": len(pos), "gen_s": round(time.time() - t0, 2)}
    for counters in (False, True):
        eng = engine.Engine(device=0, counters=counters, max_paths=8 << 20)
        t0 = time.time(); eng.set_scene(pos, nrm, uv); set_s = time.time() - t0
        eng.set_camera(engine.camera_from_settings(w, h, cam[0], cam[1], 6.0, (3.6, 3.6 * h / w), args.spp))
        for _ in range(2):
            eng.render(mode=engine.MODE_FIXED, seed=1)
        st = eng.stats()
        if counters:
            out.update(nodes_per_ray=st["nodes_visited"] / st["rays_traced"], tris_per_ray=st["tris_tested"] / st["rays_traced"])
        else:
            out.update(build_ms=st["build_ms"], set_scene_s=round(set_s, 2), n_wide_nodes=st["n_wide_nodes"], total_ms=st["total_ms"], extend_ms=st["extend_ms"],
                       shade_ms=st["shade_ms"], rays=st["rays_traced"], mrays=st["rays_traced"] / st["total_ms"] / 1e3,
                       extend_mrays=st["rays_traced"] / st["extend_ms"] / 1e3)
        eng.close()
    print(json.dumps(out), flush=True)

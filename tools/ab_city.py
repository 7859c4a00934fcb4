#!/usr/bin/env python
"""The C4 city at 32 x 32 blocks (12 M triangles, out of cache) on one GPU, 1080p, 16 spp, FIXED: per-stage ms and visits per ray, once per
value of an environment knob that the library reads at launch time:   python tools/ab_city.py [KNOB=v1,v2,...]
(profiles/r02h_ab_popcull.txt was made with VMN_POPCULL=0,1 on the build that carried the pop-time culling variant.)"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vermilion_b200 import engine, scenes
KNOB, _, vals = (sys.argv[1] if len(sys.argv) > 1 else "").partition("=")
VALUES = vals.split(",") if vals else [""]
pos, nrm, uv = scenes.city(32, 32, 23, 2024)
cam = engine.camera_from_settings(1920, 1080, (0.0, 900.0, 1950.0), (25.0, 0.0, 0.0), 6.0, (3.6, 3.6 * 1080 / 1920), 16)
for counters in (False, True):
    with engine.Engine(device=0, counters=counters, max_paths=32 << 20) as e:
        e.set_scene(pos, nrm, uv)
        e.set_camera(cam)
        for cull in VALUES:
            if KNOB:
                os.environ[KNOB] = cull
            best = None
            for i in range(3):
                e.render(mode=engine.MODE_FIXED, seed=1, max_segments=5)
                st = e.stats()
                if best is None or st["total_ms"] < best["total_ms"]:
                    best = st
            r = best["rays_traced"]
            print(json.dumps({"triangles": len(pos), "knob": KNOB, "value": cull, "counters": counters, "total_ms": round(best["total_ms"], 2), "extend_ms": round(best["extend_ms"], 2),
                              "mrays": round(r / best["total_ms"] / 1e3, 1), "nodes_per_ray": round(best["nodes_visited"] / r, 3), "tris_per_ray": round(best["tris_tested"] / r, 3)}), flush=True)

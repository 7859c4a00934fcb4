#!/usr/bin/env python
"""The C2 frame rendered as interleaved tiles on ONE GPU (tile_world = 1 and the share of rank r of `world`), per-stage device times, once per
library variant:  python tools/ab_tiles.py [--spp 16] [--world 1] default rgtileatomic
The film of tile_world = 1 is checked against the contiguous render of the same library (bit-exact)."""
import argparse, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

def child(a):
    sys.path.insert(0, ROOT)
    import numpy as np
    from vermilion_b200 import engine, scenes
    pos, nrm, uv = scenes.spheres(4, 3, 6, 150.0)
    e = engine.Engine(device=0, max_paths=128 << 20)
    e.set_scene(pos, nrm, uv)
    e.set_camera(engine.camera_from_settings(1920, 1080, (0.0, 500.0, 1900.0), (0.0, 0.0, 0.0), 6.0, (3.6, 3.6 * 1080 / 1920), a.spp))
    out = {"lib": os.path.basename(engine.LIB_PATH), "world": a.world}
    e.render(mode=engine.MODE_FIXED, seed=1, max_segments=5)
    ref = e.read_image(engine.FILM_SUM)
    best = None
    for i in range(4):
        e.render(mode=engine.MODE_FIXED, seed=1, max_segments=5, tile_size=64, tile_rank=0, tile_world=a.world)
        st = e.stats()
        if i and (best is None or st["total_ms"] < best["total_ms"]):
            best = st
    out.update({k: round(best[k], 3) for k in ("total_ms", "raygen_ms", "extend_ms", "shade_ms")})
    out["mrays"] = round(best["rays_traced"] / best["total_ms"] / 1e3, 1)
    if a.world == 1:
        out["film_equals_contiguous"] = bool(np.array_equal(ref, e.read_image(engine.FILM_SUM)))
    print(json.dumps(out))

ap = argparse.ArgumentParser()
ap.add_argument("variants", nargs="*", default=["default"])
ap.add_argument("--spp", type=int, default=16)
ap.add_argument("--world", type=int, default=1)
ap.add_argument("--child", action="store_true")
a = ap.parse_args()
if a.child:
    child(a)
else:
    for v in a.variants:
        env = dict(os.environ)
        if v != "default":
            env["VMN_LIB"] = os.path.join(ROOT, "vermilion_b200", "variants", f"libvermilion_b200_{v}.so")
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", "--spp", str(a.spp), "--world", str(a.world)], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        print(v, r.stdout.strip() or r.stderr[-600:], flush=True)

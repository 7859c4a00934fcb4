#!/usr/bin/env python
"""Two frames of the C2 bench workload at reduced spp (for ncu): python tools/c2_once.py [spp]   (one batch of spp x 1920 x 1080 paths)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vermilion_b200 import engine, scenes
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16
pos, nrm, uv = scenes.spheres(4, 3, 6, 150.0)
eng = engine.Engine(device=0, max_paths=32 << 20)
eng.set_scene(pos, nrm, uv)
eng.set_camera(engine.camera_from_settings(1920, 1080, (0.0, 500.0, 1900.0), (0.0, 0.0, 0.0), 6.0, (3.6, 3.6 * 1080 / 1920), spp))
for _ in range(2):
    eng.render(mode=engine.MODE_FIXED, seed=1, max_segments=5)
st = eng.stats()
print("ok", st["rays_traced"], st["kernel_launches"], st["extend_ms"], st["shade_ms"])

#!/usr/bin/env python
"""Two frames of the C4 city at 32 x 32 blocks (12 M triangles, out of cache), 1080p, FIXED (for ncu): python tools/city_once.py [spp] [blocks]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vermilion_b200 import engine, scenes
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
blocks = int(sys.argv[2]) if len(sys.argv) > 2 else 32
pos, nrm, uv = scenes.city(blocks, blocks, 23, 2024)
eng = engine.Engine(device=0, max_paths=32 << 20)
eng.set_scene(pos, nrm, uv)
eng.set_camera(engine.camera_from_settings(1920, 1080, (0.0, 900.0, 1950.0), (25.0, 0.0, 0.0), 6.0, (3.6, 3.6 * 1080 / 1920), spp))
for _ in range(2):
    eng.render(mode=engine.MODE_FIXED, seed=1, max_segments=5)
st = eng.stats()
print("ok", len(pos), st["rays_traced"], st["kernel_launches"], st["extend_ms"], st["shade_ms"], st["total_ms"])

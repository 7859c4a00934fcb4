#!/usr/bin/env python
"""One k_intersect launch on the C3 soup (for ncu): python tools/c3_once.py [tris] [rays]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vermilion_b200 import engine, scenes
tris = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
nrays = int(sys.argv[2]) if len(sys.argv) > 2 else 4_194_304
pos, _, _ = scenes.soup(tris, seed=1234)
rays = scenes.uniform_rays(nrays, seed=4321, lo=0.0, hi=1000.0)
eng = engine.Engine(device=0)
eng.set_scene(pos, None, None, spheres=np.zeros(0, engine.SPHERE))
d_rays = torch.from_numpy(rays).cuda(); d_hits = torch.empty((nrays, 4), device="cuda")
for _ in range(2):
    eng.intersect_device(d_rays.data_ptr(), nrays, d_hits.data_ptr(), 0)
    eng.synchronize()
print("ok", eng.stats()["n_wide_nodes"])

#!/usr/bin/env python
"""Diagnostic for the overflow-report path: run ONE traversal (argv[1] = 8 or 2) with the 3-entry-stack build and the depth check off.
VMN_LIB=vermilion_b200/variants/libvermilion_b200_tinystack.so VMN_DEBUG_NO_DEPTH_CHECK=1 timeout 60 python tools/diag_tinystack.py 8 [nrays]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vermilion_b200 import engine, scenes
which = int(sys.argv[1]); n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
level = int(sys.argv[3]) if len(sys.argv) > 3 else 3
pos, nrm, uv = scenes.spheres(2, 1, level, 150.0)
rays = scenes.uniform_rays(n, seed=7, lo=-400.0, hi=900.0)
e = engine.Engine(device=0)
e.set_scene(pos, nrm, uv)
print("commit ok, depth", e.stats()["wide_depth"], flush=True)
try:
    h = e.intersect(rays, engine.ISECT_BVH2 if which == 2 else 0)
    print("intersect ok", which, int((h["tri"] != engine.MISS).sum()), flush=True)
except engine.VmnError as x:
    print("intersect reported", which, x, flush=True)

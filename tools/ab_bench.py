#!/usr/bin/env python
"""A/B harness for kernel experiments: renders the C2 frame (reduced spp) with each library variant (csrc/Makefile VARIANT=…) in
its own subprocess and prints per-stage device times.   python tools/ab_bench.py [--spp 16] [--c3] default nocoop minb8 …"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child(args):
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch
    from vermilion_b200 import engine, scenes
    torch.cuda.set_device(0)
    out = {"lib": os.path.basename(engine.LIB_PATH)}
    if args.c3:
        pos, _, _ = scenes.soup(args.c3_tris, seed=1234)
        rays = scenes.uniform_rays(args.c3_rays, seed=4321, lo=0.0, hi=1000.0)
        eng = engine.Engine(device=0)
        eng.set_scene(pos, None, None, spheres=np.zeros(0, engine.SPHERE))
        d_rays = torch.from_numpy(rays).cuda(); d_hits = torch.empty((len(rays), 4), device="cuda")
        ms = []
        for i in range(6):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); eng.synchronize()
            eng.intersect_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr(), 0)
            eng.synchronize()
            ms.append(0)
        # device time from the library's own events is only kept by the host entry; time with wall clock around sync
        import time
        ts = []
        for i in range(5):
            eng.synchronize(); t0 = time.perf_counter()
            eng.intersect_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr(), 0)
            eng.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
        out.update(c3_ms=min(ts), c3_mrays=len(rays) / min(ts) / 1e3)
        eng.close()
    else:
        pos, nrm, uv = scenes.spheres(4, 3, 6, 150.0)
        eng = engine.Engine(device=0, max_paths=args.max_paths)
        eng.set_scene(pos, nrm, uv)
        eng.set_camera(engine.camera_from_settings(1920, 1080, (0.0, 500.0, 1900.0), (0.0, 0.0, 0.0), 6.0, (3.6, 3.6 * 1080 / 1920), args.spp))
        best = None
        for i in range(4):
            eng.render(mode=engine.MODE_FIXED, seed=1, max_segments=5)
            st = eng.stats()
            if i and (best is None or st["total_ms"] < best["total_ms"]):
                best = st
        out.update({k: round(best[k], 3) for k in ("total_ms", "raygen_ms", "extend_ms", "shade_ms", "resolve_ms")})
        out["mrays"] = round(best["rays_traced"] / best["total_ms"] / 1e3, 1)
        eng.close()
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("variants", nargs="*", default=["default"])
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--c3", action="store_true")
    ap.add_argument("--c3-tris", type=int, default=10_000_000)
    ap.add_argument("--c3-rays", type=int, default=16_777_216)
    ap.add_argument("--max-paths", type=int, default=8 << 20)
    ap.add_argument("--child", action="store_true")
    args = ap.parse_args()
    if args.child:
        return child(args)
    for v in args.variants:
        env = dict(os.environ)
        # "name@K=V,K2=V2" sets environment knobs (e.g. VMN_REFILL) and/or MAXPATHS for that run
        v, _, knobs = v.partition("@")
        mp = args.max_paths
        for kv in filter(None, knobs.split(",")):
            k, _, val = kv.partition("=")
            if k == "MAXPATHS":
                mp = int(val)
            else:
                env[k] = val
        if v != "default":
            env["VMN_LIB"] = os.path.join(ROOT, "vermilion_b200", "variants", f"libvermilion_b200_{v}.so")
        cmd = [sys.executable, os.path.abspath(__file__), "--child", "--spp", str(args.spp), "--c3-tris", str(args.c3_tris), "--c3-rays", str(args.c3_rays), "--max-paths", str(mp)] + (["--c3"] if args.c3 else [])
        r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        print(v, knobs, r.stdout.strip() or r.stderr[-500:], flush=True)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""N1 row: BruteForceTracer::Render (integrators.cpp:9-186) on the C2 scene at 1080p, 128 spp (the reference's
default raysPerPixel; its ::abs(int) early-out stops almost every pixel at the third sample), VMN_MODE_BRUTEFORCE through the C ABI
against the REAL reference Render on all host cores (oracle/_ref/libvermref_ofast.so).   python tools/bench_bruteforce.py [--shadow]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shadow", action="store_true")
    ap.add_argument("--res", default="1920x1080")
    ap.add_argument("--spp", type=int, default=128)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import numpy as np
    import torch
    from vermilion_b200 import engine, scenes
    import checker
    w, h = (int(x) for x in args.res.split("x"))
    pos, nrm, uv = scenes.spheres(4, 3, 6, 150.0)
    cam_pos, cam_rot = (-900.0, 700.0, 1800.0), (-10.0, 25.0, 0.0)  # looks across the sphere grid: about a third of the primary rays hit triangles
    cam = engine.camera_from_settings(w, h, cam_pos, cam_rot, 6.0, (3.6, 3.6 * h / w), args.spp)
    out = {"workload": f"BruteForceTracer, C2 scene (983040 triangles), {w}x{h}, {args.spp} spp requested, shadow rays {'on' if args.shadow else 'off'}"}
    with engine.Engine(device=0, max_paths=8 << 20) as eng:
        eng.set_scene(pos, nrm, uv)
        eng.set_camera(cam)
        pimg = torch.empty((w * h, 5), dtype=torch.float32).pin_memory().numpy()
        best = None
        for i in range(args.steps + 2):
            t0 = time.perf_counter()
            eng.render(mode=engine.MODE_BRUTEFORCE, seed=1, shadow=args.shadow)
            img = eng.read_image(out=pimg)
            wall = (time.perf_counter() - t0) * 1e3
            st = eng.stats()
            st["e2e_ms"] = wall
            if i >= 2 and (best is None or st["total_ms"] < best["total_ms"]):
                best = st
        out.update(gpu_ms=best["total_ms"], gpu_e2e_ms=best["e2e_ms"], rays=best["rays_traced"], samples=best["paths_traced"],
                   gpu_mrays=best["rays_traced"] / best["total_ms"] / 1e3, gpu_e2e_mrays=best["rays_traced"] / best["e2e_ms"] / 1e3,
                   stages_ms={k: round(best[k], 3) for k in ("raygen_ms", "extend_ms", "shade_ms")}, launches=best["kernel_launches"],
                   triangle_pixels=float((img[:, 4] < 3000).mean()), mean_rgb=[float(x) for x in np.nan_to_num(img[:, :3]).mean(0)])
    if not args.no_cpu and checker.Ref.available("ofast") and not args.shadow:
        cs = checker.camera_settings(w, h, cam_pos, cam_rot, 6.0, (3.6, 3.6 * h / w), args.spp)
        ref = checker.Ref(pos, nrm, uv, variant="ofast")
        ref.render_bruteforce(cs, seed=1, threads=0)
        rimg, secs = ref.render_bruteforce(cs, seed=1, threads=0)
        out.update(cpu_ms=secs * 1e3, cpu_cores=checker.Oracle.max_threads(), cpu_mrays=out["rays"] / secs / 1e6,
                   rmse_vs_reference=float(np.sqrt(np.nanmean((np.nan_to_num(img[:, :3]) - np.nan_to_num(rimg[:, :3])) ** 2))),
                   cpu_note="the real BruteForceTracer::Render (-Ofast as shipped), all host threads racing on its shared mt19937 as the reference does; "
                            "rays counted by the GPU run (the reference has no counter)")
    print(json.dumps(out))


if __name__ == "__main__":
    main()

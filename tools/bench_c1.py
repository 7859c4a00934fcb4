#!/usr/bin/env python
"""C1 (BASELINE.json configs[0]): Cornell box (30 triangles) 512x512, 16 spp, the reference's own estimator (VMN_MODE_COMPAT: adaptive
early-out, roulette) — the GPU engine through the C ABI with host buffers against the REAL PathTracer::Render on all host cores
(oracle/_ref/libvermref_ofast.so, -Ofast as shipped), with the image RMSE between the two (different RNG streams).
  python tools/bench_c1.py [--spp 16] [--res 512x512]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--res", default="512x512")
    args = ap.parse_args()
    import numpy as np
    from vermilion_b200 import engine, scenes
    import checker
    w, h = (int(x) for x in args.res.split("x"))
    pos, nrm, uv = scenes.cornell()
    cam_pos, cam_rot = (0.0, 275.0, 1100.0), (0.0, 0.0, 0.0)
    cam = engine.camera_from_settings(w, h, cam_pos, cam_rot, 6.0, (3.6, 3.6 * h / w), args.spp)
    out = {"workload": f"C1: Cornell box ({len(pos)} triangles) {w}x{h}, {args.spp} spp, COMPAT (the reference estimator)"}
    with engine.Engine(device=0) as eng:
        best = None
        for i in range(5):
            t0 = time.perf_counter()
            eng.set_scene(pos, nrm, uv)
            eng.set_camera(cam)
            eng.render(mode=engine.MODE_COMPAT, seed=1 + i)
            img = eng.read_image()
            ms = (time.perf_counter() - t0) * 1e3
            st = eng.stats()
            if i and (best is None or ms < best[0]):
                best = (ms, st)
        ms, st = best
        out.update(gpu_e2e_ms=ms, gpu_render_ms=st["total_ms"], rays=st["rays_traced"], samples=st["paths_traced"], launches=st["kernel_launches"],
                   gpu_e2e_mrays=st["rays_traced"] / ms / 1e3, mean_samples_per_pixel=float(img[:, 4].mean()))
    if checker.Ref.available("ofast"):
        cs = checker.camera_settings(w, h, cam_pos, cam_rot, 6.0, (3.6, 3.6 * h / w), args.spp)
        ref = checker.Ref(pos, nrm, uv, variant="ofast")
        ref.render_pathtracer(cs)
        rimg, secs = ref.render_pathtracer(cs)
        out.update(cpu_ms=secs * 1e3, cpu_cores=checker.Oracle.max_threads(), cpu_mean_samples_per_pixel=float(rimg[:, 4].mean()),
                   speedup_e2e=secs * 1e3 / ms, rmse_vs_reference=float(np.sqrt(np.mean((img[:, :3] - rimg[:, :3]) ** 2))),
                   cpu_note="the real PathTracer::Render (-Ofast as shipped, OpenMP over all host threads, std::random_device seeds)")
        a, _ = ref.render_pathtracer(cs)
        out["rmse_reference_vs_reference"] = float(np.sqrt(np.mean((a[:, :3] - rimg[:, :3]) ** 2)))
    print(json.dumps(out))


if __name__ == "__main__":
    main()

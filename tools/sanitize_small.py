#!/usr/bin/env python
"""Small-scene pass over every kernel of the library for compute-sanitizer: run as

  compute-sanitizer --tool memcheck  python tools/sanitize_small.py
  compute-sanitizer --tool racecheck python tools/sanitize_small.py
  compute-sanitizer --tool synccheck python tools/sanitize_small.py

Touches build (k_prim_bounds .. k_collapse), k_intersect (BVH8 + BVH2, counting and not; also through vmn_intersect_static), k_occluded, k_raycast, vmn_radiance (both
estimators), k_raygen / k_extend / k_shade / k_accumulate / k_resolve (FIXED, COMPAT, tiles, accumulate), the BRUTEFORCE kernels (with and
without shadow rays, textured), k_pack_rgba8 and the bandwidth probe. Results are compared with the CPU oracle so that a sanitizer-clean
run is also a correct one. Sizes are tiny: the sanitizer slows kernels down 10-100x."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]


def main():
    import checker
    import common
    from vermilion_b200 import engine, scenes
    pos, nrm, uv = scenes.spheres(2, 1, 2, 150.0)  # 640 triangles
    soup, _, _ = scenes.soup(3000, seed=1234)
    box = (np.array([-1900, 10, -1900], np.float32), np.array([1900, 990, 1900], np.float32))
    rays = common.aimed_rays(pos, 3000, 1, spread=0.05, origin_box=box)
    srays = scenes.uniform_rays(3000, seed=4321, lo=0.0, hi=1000.0)
    rng = np.random.default_rng(3)
    tex = rng.random((8, 8, 3), dtype=np.float32)
    tex1 = rng.random((16, 8, 4), dtype=np.float32)
    orc = checker.Oracle(pos, nrm, uv)
    for counters in (False, True):
        with engine.Engine(device=0, counters=counters, max_paths=1 << 14) as e:
            e.set_scene(soup, None, None, spheres=np.zeros(0, engine.SPHERE))
            a, b = e.intersect(srays), e.intersect(srays, engine.ISECT_BVH2)
            assert (a["tri"] == b["tri"]).mean() > 0.999
            e.set_scene(pos, nrm, uv)
            h = e.intersect(rays)
            c = common.classify_hits(h, orc.intersect(rays))
            assert c["bad"] == 0 and c["t_exact"] == c["hits"], c
            for _ in range(2):  # the static-scratch entry point (vmn_intersect_static), twice over the same array
                assert np.array_equal(e.intersect_static(rays).view(np.uint8), h.view(np.uint8))
            assert np.array_equal(e.occluded(rays), orc.occluded(rays))
            rc = e.raycast(rays)
            assert (rc["flags"] & 1).mean() > 0.9
            for mode in (engine.MODE_COMPAT, engine.MODE_FIXED):
                out, cnt = e.radiance(rays[:1000], 5, mode)
                assert np.isfinite(out[:, 3]).all() and cnt.max() >= 1
            cam = engine.camera_from_settings(48, 32, (0.0, 500.0, 1900.0), (0.0, 0.0, 0.0), 6.0, (3.6, 2.4), 8)
            e.set_camera(cam)
            ocam = checker.OrcCamera(cam.width, cam.height, cam.position, cam.rotation, cam.dist_to_film, cam.sensor_x, cam.sensor_y, cam.spp)
            for mode, omode in ((engine.MODE_FIXED, checker.MODE_FIXED), (engine.MODE_COMPAT, checker.MODE_COMPAT)):
                e.render(mode=mode, seed=3)
                img = e.read_image()
                oimg, _ = orc.render(ocam, mode=omode, rng=checker.RNG_PHILOX, seed=3)
                assert (img.view(np.uint32) == oimg.view(np.uint32)).all(1).mean() > 0.97
            full = e.read_image(engine.FILM_SUM)
            e.render(mode=engine.MODE_FIXED, seed=3, tile_size=16, tile_rank=1, tile_world=2)
            e.render(mode=engine.MODE_FIXED, seed=3, sample_begin=0, sample_end=4)
            e.render(mode=engine.MODE_FIXED, seed=3, sample_begin=4, sample_end=8, accumulate=True)
            e.read_image_rgba8()
            e.set_texture(0, tex)
            e.set_texture(1, tex1)
            e.render(mode=engine.MODE_FIXED, seed=3)
            for shadow in (False, True):
                e.render(mode=engine.MODE_BRUTEFORCE, seed=3, shadow=shadow)
                assert np.isfinite(e.read_image()[:, :3]).all()
            e.probe_read_bandwidth(1 << 20, 1)
            e.synchronize()
            del full
    print("sanitize_small: all kernels ran, results match the oracle")


if __name__ == "__main__":
    main()

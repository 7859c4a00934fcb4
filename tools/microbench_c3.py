#!/usr/bin/env python
"""C3 — the traversal-roofline row (BASELINE.json configs[2]): N random rays against an M-triangle soup, BVH-only closest hit
through vmn_intersect_device (rays and hits resident in HBM). Prints one JSON line.

  python tools/microbench_c3.py [--tris 10000000] [--rays 16777216] [--reps 5] [--check 200000] [--bvh2]

Parity at full size: (1) a --check sized sample of the same rays against the CPU oracle (bit-exact t/u/v, ids exact up to
reported ties); (2) all rays: the BVH8 traversal against the independent binary-LBVH traversal of the same library
(different tree, same closest hits — a size-independent property)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=10_000_000)
    ap.add_argument("--rays", type=int, default=16_777_216)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--check", type=int, default=200_000)
    ap.add_argument("--no-bvh2", action="store_true")
    ap.add_argument("--sort", action="store_true", help="also time Morton-sorted rays (coherent order)")
    args = ap.parse_args()
    import torch
    from vermilion_b200 import engine, scenes
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    t0 = time.time()
    pos, nrm, uv = scenes.soup(args.tris, seed=1234)
    rays = scenes.uniform_rays(args.rays, seed=4321, lo=0.0, hi=1000.0)
    gen_s = time.time() - t0
    out = {"workload": f"C3: {args.rays} uniform random rays vs {args.tris}-triangle soup in [0,1000]^3, BVH-only closest hit", "gen_s": gen_s}

    def run(counters):
        eng = engine.Engine(device=0, counters=counters, stream=stream.cuda_stream)
        eng.set_scene(pos, None, None, spheres=np.zeros(0, engine.SPHERE))
        return eng

    eng = run(False)
    st = eng.stats()
    out.update(build_ms=st["build_ms"], n_wide_nodes=st["n_wide_nodes"], node_bytes=st["node_bytes"], tri_bytes=st["tri_bytes"])
    d_rays = torch.from_numpy(rays).cuda()
    d_hits = torch.empty((args.rays, 4), dtype=torch.float32, device="cuda")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")

    def timed(flags, d_r=d_rays):
        ms = []
        for i in range(args.reps + 2):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            eng.intersect_device(d_r.data_ptr(), args.rays, d_hits.data_ptr(), flags)
            b.record(stream)
            b.synchronize()
            if i >= 2:
                ms.append(a.elapsed_time(b))
        return float(np.mean(ms)), float(np.min(ms))

    ms8, ms8_min = timed(0)
    h8 = d_hits.cpu().numpy().copy().view(engine.HIT).reshape(-1)
    out.update(ms_bvh8=ms8, ms_bvh8_min=ms8_min, mrays_bvh8=args.rays / ms8 / 1e3, hit_rate=float((h8["tri"] != engine.MISS).mean()))
    if not args.no_bvh2:
        ms2, _ = timed(engine.ISECT_BVH2)
        h2 = d_hits.cpu().numpy().copy().view(engine.HIT).reshape(-1)
        diff = h8["tri"] != h2["tri"]
        same_t = np.abs(h8["t"][diff] - h2["t"][diff]) <= 4e-7 * np.maximum(1.0, np.abs(h2["t"][diff]))
        out.update(ms_bvh2=ms2, mrays_bvh2=args.rays / ms2 / 1e3, bvh8_vs_bvh2_id_mismatch=int(diff.sum()), bvh8_vs_bvh2_ties=int(same_t.sum()),
                   bvh8_vs_bvh2_t_bitexact=bool(np.array_equal(h8["t"][~diff], h2["t"][~diff])))
    if args.sort:
        o = rays[:, :3]
        q = np.clip((o / 1000.0 * 1024).astype(np.int64), 0, 1023)
        def spread(x):
            x = (x | (x << 16)) & 0x030000FF; x = (x | (x << 8)) & 0x0300F00F; x = (x | (x << 4)) & 0x030C30C3; x = (x | (x << 2)) & 0x09249249
            return x
        key = (spread(q[:, 0]) << 2) | (spread(q[:, 1]) << 1) | spread(q[:, 2])
        d_sorted = torch.from_numpy(rays[np.argsort(key, kind="stable")]).cuda()
        mss, _ = timed(0, d_sorted)
        out.update(ms_bvh8_origin_sorted=mss, mrays_bvh8_origin_sorted=args.rays / mss / 1e3)
    eng.close()

    ce = run(True)
    # visits per ray from the counting build on a 2 M sample of the same rays (host entry point)
    ns = min(args.rays, 2_000_000)
    ce.intersect(rays[:ns])
    cst = ce.stats()
    ce.close()
    npr, tpr = cst["nodes_visited"] / ns, cst["tris_tested"] / ns
    b_ray = 32 + 16 + npr * cst["node_bytes"] + tpr * cst["tri_bytes"]
    peak = 6549.8
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    ach = args.rays / (ms8 / 1e3) * b_ray / 1e9
    out["roofline"] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "bytes_per_ray": b_ray,
                       "nodes_per_ray": npr, "tris_per_ray": tpr}

    if args.check:
        import checker
        import common
        n = min(args.check, args.rays)
        t0 = time.time()
        orc = checker.Oracle(pos)
        ref = orc.intersect(rays[:n].view(checker.RAY).reshape(-1))
        c = common.classify_hits(h8[:n], ref)
        out["oracle_check"] = dict(c, seconds=time.time() - t0)
        rn, rt = orc.intersect_stats(rays[:n].view(checker.RAY).reshape(-1))
        out["reference_tree_visits"] = {"nodes_per_ray": rn / n, "tris_per_ray": rt / n, "note": "binary midpoint-split BVH of bvh.cpp:179-279, 48-byte nodes"}
        t0 = time.time()
        a, b = orc.intersect(rays[:n].view(checker.RAY).reshape(-1)), None
        cpu_s = time.time() - t0
        out["cpu_port"] = {"mrays": n / cpu_s / 1e6, "cores": checker.Oracle.max_threads(), "sample": f"{n} of the same rays, oracle port, tree built once"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()

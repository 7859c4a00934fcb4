#!/usr/bin/env python
"""C3 — the traversal-roofline row (BASELINE.json configs[2]): N random rays against an M-triangle soup, BVH-only closest hit
through vmn_intersect_device (rays and hits resident in HBM). Prints one JSON line; bench.py embeds the same record as
`roofline_c3` in its N = 1 line (c3_record below).

  python tools/microbench_c3.py [--tris 10000000] [--rays 16777216] [--reps 5] [--check 200000] [--no-bvh2]

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

C3_TRIS, C3_RAYS = 10_000_000, 16_777_216


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def c3_record(stream, tris=C3_TRIS, n_rays=C3_RAYS, reps=5, check=200_000, bvh2=True, sort=False, device=0, flush=None):
    """Runs C3 on `device` (torch CUDA stream `stream`) and returns the record. Rays: std::mt19937-seeded uniform origins in the cube and
    directions on the sphere (scenes.uniform_rays, seed 4321) against scenes.soup(tris, seed 1234). Timing: CUDA events on `stream`
    around vmn_intersect_device, L2 flushed (512 MiB write) before every rep, 2 untimed reps first."""
    import torch
    from vermilion_b200 import engine, scenes
    t0 = time.time()
    pos, _, _ = scenes.soup(tris, seed=1234)
    rays = scenes.uniform_rays(n_rays, seed=4321, lo=0.0, hi=1000.0)
    out = {"workload": f"C3: {n_rays} uniform random rays vs {tris}-triangle soup in [0,1000]^3, BVH-only closest hit (vmn_intersect_device)",
           "gen_s": time.time() - t0}

    def run(counters):
        eng = engine.Engine(device=device, counters=counters, stream=stream.cuda_stream)
        eng.set_scene(pos, None, None, spheres=np.zeros(0, engine.SPHERE))
        return eng

    eng = run(False)
    st = eng.stats()
    out.update(build_ms=st["build_ms"], n_wide_nodes=st["n_wide_nodes"], wide_depth=st["wide_depth"], node_bytes=st["node_bytes"], tri_bytes=st["tri_bytes"])
    d_rays = torch.from_numpy(rays).cuda()
    d_hits = torch.empty((n_rays, 4), dtype=torch.float32, device="cuda")
    if flush is None:
        flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")

    def timed(flags, d_r=d_rays):
        ms = []
        for i in range(reps + 2):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            eng.intersect_device(d_r.data_ptr(), n_rays, d_hits.data_ptr(), flags)
            b.record(stream)
            b.synchronize()
            if i >= 2:
                ms.append(a.elapsed_time(b))
        eng.synchronize()  # raises if a traversal stack overflowed
        return float(np.mean(ms)), float(np.min(ms))

    ms8, ms8_min = timed(0)
    h8 = d_hits.cpu().numpy().copy().view(engine.HIT).reshape(-1)
    out.update(ms_bvh8=ms8, ms_bvh8_min=ms8_min, mrays_bvh8=n_rays / ms8 / 1e3, hit_rate=float((h8["tri"] != engine.MISS).mean()), reps=reps,
               l2=f"{flush.numel() >> 20} MiB L2 flush before every rep")
    if bvh2:
        ms2, _ = timed(engine.ISECT_BVH2)
        h2 = d_hits.cpu().numpy().copy().view(engine.HIT).reshape(-1)
        diff = h8["tri"] != h2["tri"]
        same_t = np.abs(h8["t"][diff] - h2["t"][diff]) <= 4e-7 * np.maximum(1.0, np.abs(h2["t"][diff]))
        out.update(ms_bvh2=ms2, mrays_bvh2=n_rays / ms2 / 1e3, bvh8_vs_bvh2_id_mismatch=int(diff.sum()), bvh8_vs_bvh2_ties=int(same_t.sum()),
                   bvh8_vs_bvh2_t_bitexact=bool(np.array_equal(h8["t"][~diff], h2["t"][~diff])))
    if sort:
        o = rays[:, :3]
        q = np.clip((o / 1000.0 * 1024).astype(np.int64), 0, 1023)

        def spread(x):
            x = (x | (x << 16)) & 0x030000FF; x = (x | (x << 8)) & 0x0300F00F; x = (x | (x << 4)) & 0x030C30C3; x = (x | (x << 2)) & 0x09249249
            return x
        key = (spread(q[:, 0]) << 2) | (spread(q[:, 1]) << 1) | spread(q[:, 2])
        d_sorted = torch.from_numpy(rays[np.argsort(key, kind="stable")]).cuda()
        mss, _ = timed(0, d_sorted)
        out.update(ms_bvh8_origin_sorted=mss, mrays_bvh8_origin_sorted=n_rays / mss / 1e3)
    eng.close()

    ce = run(True)
    # visits per ray from the counting build on a 2 M sample of the same rays (host entry point)
    ns = min(n_rays, 2_000_000)
    ce.intersect(rays[:ns])
    cst = ce.stats()
    ce.close()
    npr, tpr = cst["nodes_visited"] / ns, cst["tris_tested"] / ns
    b_ray = 32 + 16 + npr * cst["node_bytes"] + tpr * cst["tri_bytes"]
    peak, peak_src = hbm_peak()
    ach = n_rays / (ms8 / 1e3) * b_ray / 1e9
    out["roofline"] = {"bound": "hbm", "kernel": "k_intersect<true,false>", "achieved": ach, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                       "frac": ach / peak, "bytes_per_ray": b_ray, "nodes_per_ray": npr, "tris_per_ray": tpr,
                       "note": "B_ray = 32 (ray) + 16 (hit) + nodes/ray x 80 + tris/ray x 48 (SURVEY 8d); visits counted by the VMN_CFG_COUNTERS build on "
                               f"the first {ns} of the same rays"}
    if check:
        import checker
        import common
        n = min(check, n_rays)
        t0 = time.time()
        orc = checker.Oracle(pos)
        ref = orc.intersect(rays[:n].view(checker.RAY).reshape(-1))
        c = common.classify_hits(h8[:n], ref)
        out["oracle_check"] = dict(c, seconds=time.time() - t0,
                                   bit_exact=bool(c["bad"] == 0 and c["t_exact"] == c["hits"] and c["uv_err_max"] == 0.0))
        rn, rt = orc.intersect_stats(rays[:n].view(checker.RAY).reshape(-1))
        out["reference_tree_visits"] = {"nodes_per_ray": rn / n, "tris_per_ray": rt / n, "note": "binary midpoint-split BVH of bvh.cpp:179-279, 48-byte nodes"}
        t0 = time.time()
        orc.intersect(rays[:n].view(checker.RAY).reshape(-1))
        cpu_s = time.time() - t0
        out["cpu_port"] = {"mrays": n / cpu_s / 1e6, "cores": checker.Oracle.max_threads(), "sample": f"{n} of the same rays, oracle port, tree built once"}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=C3_TRIS)
    ap.add_argument("--rays", type=int, default=C3_RAYS)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--check", type=int, default=200_000)
    ap.add_argument("--no-bvh2", action="store_true")
    ap.add_argument("--sort", action="store_true", help="also time Morton-sorted rays (coherent order)")
    args = ap.parse_args()
    import torch
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    print(json.dumps(c3_record(stream, args.tris, args.rays, args.reps, args.check, not args.no_bvh2, args.sort)))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Upper bound of what binning secondary rays can buy, measured before building it into the wavefront:
bounce rays of a real frame are generated the way k_shade makes them (primary hit -> cosine-weighted direction about the normal, origin on
the surface), kept in QUEUE order (= pixel order, as the compacted queues are), and k_intersect is timed on (a) that order, (b) the same
rays binned by direction octant x coarse origin cell (8 x 16^3 bins, stable counting sort = what a binned append would produce), (c)
fully Morton-sorted by origin within each octant. Two generations of bounces (bounce 1 from camera hits, bounce 2 from bounce-1 hits).

  python tools/ab_coherence.py [--scene c2|city:G] [--res 1920x1080] [--reps 5]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]


def cosine_dirs(n, rng):
    r1 = 2 * np.pi * rng.random(len(n)); r2 = rng.random(len(n)); r2s = np.sqrt(r2)
    w = n
    a = np.where((np.abs(w[:, 0]) > 0.1)[:, None], np.array([0, 1, 0], np.float64), np.array([1, 0, 0], np.float64))
    u = np.cross(a, w); u /= np.linalg.norm(u, axis=1, keepdims=True)
    v = np.cross(w, u)
    d = u * (np.cos(r1) * r2s)[:, None] + v * (np.sin(r1) * r2s)[:, None] + w * np.sqrt(1 - r2)[:, None]
    return d / np.linalg.norm(d, axis=1, keepdims=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="c2")
    ap.add_argument("--res", default="1920x1080")
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    import torch
    from vermilion_b200 import engine, scenes
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
    w, h = (int(x) for x in args.res.split("x"))
    if args.scene == "c2":
        pos, nrm, uv = scenes.spheres(4, 3, 6, 150.0); cam_pos, cam_rot = (0.0, 500.0, 1900.0), (0.0, 0.0, 0.0)
    else:
        g = int(args.scene.split(":")[1]); pos, nrm, uv = scenes.city(g, g, 23, 2024); cam_pos, cam_rot = (0.0, 900.0, 1950.0), (25.0, 0.0, 0.0)
    eng = engine.Engine(device=0, stream=stream.cuda_stream)
    eng.set_scene(pos, nrm, uv)
    eng.set_camera(engine.camera_from_settings(w, h, cam_pos, cam_rot, 6.0, (3.6, 3.6 * h / w), 4))
    rng = np.random.default_rng(1)
    n = w * h
    px = np.arange(n, dtype=np.uint64)
    rays = eng.camera_rays(px, np.zeros(n, np.uint32), np.zeros(n, np.uint32), (rng.random(n) * 0.5).astype(np.float32), (rng.random(n) * 0.5).astype(np.float32))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    d_hits = torch.empty((n, 4), dtype=torch.float32, device="cuda")

    def timed(r):
        d = torch.from_numpy(np.ascontiguousarray(r).view(np.float32).reshape(-1, 8)).cuda()
        ms = []
        for i in range(args.reps + 2):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream); eng.intersect_device(d.data_ptr(), len(r), d_hits.data_ptr(), 0); b.record(stream); b.synchronize()
            if i >= 2:
                ms.append(a.elapsed_time(b))
        return float(np.mean(ms))

    def keys(r, cells):
        o, d = r["o"].astype(np.float64), r["d"]
        lo, hi = o.min(0), o.max(0)
        q = np.clip(((o - lo) / np.maximum(hi - lo, 1e-9) * cells).astype(np.int64), 0, cells - 1)

        def spread(x):
            x = (x | (x << 16)) & 0x030000FF; x = (x | (x << 8)) & 0x0300F00F; x = (x | (x << 4)) & 0x030C30C3; x = (x | (x << 2)) & 0x09249249
            return x
        m = (spread(q[:, 0]) << 2) | (spread(q[:, 1]) << 1) | spread(q[:, 2])
        octant = ((d[:, 0] >= 0).astype(np.int64) << 2) | ((d[:, 1] >= 0).astype(np.int64) << 1) | (d[:, 2] >= 0).astype(np.int64)
        return (octant << 40) | m

    out = {"scene": args.scene, "triangles": int(len(pos)), "rays_per_set": n, "sets": {}}
    cur = rays
    for gen in (0, 1, 2):
        name = ["primary", "bounce1", "bounce2"][gen]
        rec = {"queue_order_ms": timed(cur)}
        if gen:
            rec["binned_octant_x_16cubed_ms"] = timed(cur[np.argsort(keys(cur, 16), kind="stable")])
            rec["sorted_octant_x_morton1024_ms"] = timed(cur[np.argsort(keys(cur, 1024), kind="stable")])
            rec["gain_binned"] = rec["queue_order_ms"] / rec["binned_octant_x_16cubed_ms"]
            rec["gain_sorted"] = rec["queue_order_ms"] / rec["sorted_octant_x_morton1024_ms"]
        out["sets"][name] = rec
        if gen == 2:
            break
        rc = eng.raycast(cur)  # full RayCast: triangles and the room's spheres, as the wavefront's hit
        ok = (rc["flags"] & 1) != 0
        nrm_ = rc["N"][ok].astype(np.float64)
        din = cur["d"][ok].astype(np.float64)
        nrm_ = np.where((np.einsum("ij,ij->i", nrm_, din) < 0)[:, None], nrm_, -nrm_)
        good = np.isfinite(nrm_).all(1) & (np.linalg.norm(nrm_, axis=1) > 0.5)
        P = rc["P"][ok][good].astype(np.float64) - din[good] * 0.001
        nd = cosine_dirs(nrm_[good], rng)
        nxt = np.zeros(len(P), engine.RAY)
        nxt["o"] = P.astype(np.float32); nxt["d"] = nd.astype(np.float32)
        cur = nxt
    print(json.dumps(out))


if __name__ == "__main__":
    main()

"""Shared helpers for the parity tests: ray sets aimed at geometry and mismatch classification."""
import numpy as np

import checker


def aimed_rays(pos, n, seed, spread=0.0, origin_box=None):
    """n rays from random origins towards random points ON random triangles (hit rate ~1), optional angular spread."""
    rng = np.random.default_rng(seed)
    pos = np.asarray(pos, np.float32).reshape(-1, 3, 3)
    lo, hi = pos.reshape(-1, 3).min(0), pos.reshape(-1, 3).max(0)
    if origin_box is None:
        ext = np.maximum(hi - lo, 1.0)
        origin_box = (lo - 0.5 * ext, hi + 0.5 * ext)
    o = rng.uniform(origin_box[0], origin_box[1], (n, 3)).astype(np.float32)
    t = rng.integers(0, len(pos), n)
    b = rng.dirichlet((1, 1, 1), n).astype(np.float32)
    tgt = (pos[t] * b[:, :, None]).sum(1)
    d = tgt - o
    if spread:
        d = d + rng.normal(0, spread, d.shape) * np.linalg.norm(d, axis=1, keepdims=True)
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    return checker.make_rays(o, d)


def classify_hits(gpu, ref, tol=1e-5):
    """Compare closest-hit records. Returns dict with counts; `bad` must be 0.
    tie  : ids differ but t agrees within 4 ulp-ish relative (equal-distance hit on a shared edge/vertex or coincident faces)
    edge : one side hit, the other missed or hit further away, and the nearer hit lies within 1e-4 of the triangle's border
           (barycentric u, v or 1-u-v) — the box/edge rounding cases north_star allows to be excluded, reported
    """
    n = len(ref)
    same_id = gpu["tri"] == ref["tri"]
    both_hit = same_id & (ref["tri"] != checker.MISS)
    t_exact = int(np.count_nonzero(gpu["t"][both_hit] == ref["t"][both_hit]))
    rel = np.abs(gpu["t"][both_hit] - ref["t"][both_hit]) / np.maximum(1.0, np.abs(ref["t"][both_hit]))
    uv_err = np.maximum(np.abs(gpu["u"][both_hit] - ref["u"][both_hit]), np.abs(gpu["v"][both_hit] - ref["v"][both_hit]))
    diff = np.nonzero(~same_id)[0]
    tie = edge = bad = 0
    for i in diff:
        g, r = gpu[i], ref[i]
        if g["tri"] != checker.MISS and r["tri"] != checker.MISS and abs(float(g["t"]) - float(r["t"])) <= 4e-7 * max(1.0, abs(float(r["t"]))):
            tie += 1
            continue
        near = g if (g["tri"] != checker.MISS and (r["tri"] == checker.MISS or g["t"] < r["t"])) else r
        u, v = float(near["u"]), float(near["v"])
        if min(u, v, 1.0 - u - v) < 1e-4:
            edge += 1
        else:
            bad += 1
    return {"n": n, "same_id": int(same_id.sum()), "hits": int(both_hit.sum()), "t_exact": t_exact,
            "t_rel_max": float(rel.max()) if len(rel) else 0.0, "uv_err_max": float(uv_err.max()) if len(uv_err) else 0.0,
            "tie": tie, "edge": edge, "bad": bad}

"""GPU test (-m gpu), collected last on purpose: vmn_intersect_static — the closest-hit query whose device scratch is the library's
file-scope `static __device__` array (no cudaMalloc inside the call; include/vermilion_b200.h) — against vmn_intersect, the existing
entry point, through the C ABI. Bar: bit-exact records (the tolerance of the other ray-query tests: integer ids exact, t/u/v equal
bit for bit), on two calls in a row that share the array, then the size limit and the edge cases. Passes alone in a fresh process
and after the rest of the suite (it makes its own context and scene)."""
import ctypes as C

import numpy as np
import pytest

import common
from vermilion_b200 import engine, scenes

pytestmark = pytest.mark.gpu


def _same(a, b):
    return np.array_equal(a.view(np.uint8), b.view(np.uint8))


def _rays(pos, n, seed):
    a = scenes.uniform_rays(n // 2, seed=seed, lo=0.0, hi=1000.0).view(engine.RAY).reshape(-1)
    b = common.aimed_rays(pos, n - n // 2, seed + 1)
    return np.concatenate([a, b])


def test_static_scratch_entry_point_matches_vmn_intersect_twice_in_a_row():
    import torch
    pos, nrm, uv = scenes.soup(100_000, seed=1234)
    with engine.Engine(device=0, counters=True) as eng:
        eng.set_scene(pos, nrm, uv)
        first = _rays(pos, engine.STATIC_MAX_RAYS, 21)       # the whole array
        second = _rays(pos, 20_001, 33)                      # a shorter, ragged batch over the first call's leftovers
        want1, want2 = eng.intersect(first), eng.intersect(second)
        st_ref = eng.stats()
        assert (want1["tri"] != engine.MISS).mean() > 0.3 and (want2["tri"] != engine.MISS).mean() > 0.3

        torch.cuda.mem_get_info(0)                           # (torch's own start-up is not the call's doing)
        got1 = eng.intersect_static(first)
        free_before = torch.cuda.mem_get_info(0)[0]
        got2 = eng.intersect_static(second)
        st = eng.stats()
        free_after = torch.cuda.mem_get_info(0)[0]
        assert _same(got1, want1)
        assert _same(got2, want2)
        assert free_after >= free_before, "the call took device memory"
        # the visit counters live in the static array as well and are cleared per call
        assert st["rays_traced"] == len(second) and st["kernel_launches"] == 1
        # (a lane's deferred triangle tests depend on which rays share its warp, so the totals may move a little from run to run)
        assert st["nodes_visited"] > 0 and st["tris_tested"] > 0
        assert abs(st["nodes_visited"] - st_ref["nodes_visited"]) <= 0.05 * st_ref["nodes_visited"]
        assert abs(st["tris_tested"] - st_ref["tris_tested"]) <= 0.05 * st_ref["tris_tested"]

        # binary-tree cross-check kernel through the same array, and a third call after it
        assert _same(eng.intersect_static(second, engine.ISECT_BVH2), eng.intersect(second, engine.ISECT_BVH2))
        assert _same(eng.intersect_static(first), want1)

        # edge cases: empty batch, one ray, the limit
        assert len(eng.intersect_static(first[:0])) == 0
        assert _same(eng.intersect_static(first[:1]), want1[:1])
        too_many = np.concatenate([first, first[:1]])
        with pytest.raises(engine.VmnError, match="VMN_STATIC_MAX_RAYS = 65536"):
            eng.intersect_static(too_many)
        out = np.empty(1, engine.HIT)
        assert eng.L.vmn_intersect_static(eng.h, None, 1, out.ctypes.data_as(C.c_void_p), 0) == 1  # VMN_ERR_INVALID
        assert _same(eng.intersect_static(second), want2)    # still healthy after the rejected calls


def test_static_scratch_needs_a_committed_scene():
    with engine.Engine(device=0) as eng:
        rays = np.zeros(4, engine.RAY)
        with pytest.raises(engine.VmnError, match="not committed"):
            eng.intersect_static(rays)


def test_two_contexts_share_the_array_when_serialised():
    """The array belongs to the device, not to a context: a second context (other scene) used after the first sees no leftovers."""
    pos_a, nrm_a, uv_a = scenes.soup(20_000, seed=7)
    pos_b, nrm_b, uv_b = scenes.soup(5_000, seed=8)
    with engine.Engine(device=0) as a, engine.Engine(device=0) as b:
        a.set_scene(pos_a, nrm_a, uv_a)
        b.set_scene(pos_b, nrm_b, uv_b)
        ra = _rays(pos_a, 30_000, 5)
        rb = _rays(pos_b, 10_001, 9)
        for _ in range(2):
            assert _same(a.intersect_static(ra), a.intersect(ra))
            assert _same(b.intersect_static(rb), b.intersect(rb))

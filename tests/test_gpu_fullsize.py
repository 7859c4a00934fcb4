"""GPU tests (-m gpu) at BASELINE.json's full C2 size (983 040 triangles, 1920x1080), where the CPU oracle cannot render the frame:
size-independent properties of the domain plus an oracle-checked sample.

  * two independent traversals (compressed BVH8 vs the binary LBVH) agree on 2 M rays: ids equal except reported ties / border cases,
    t and barycentrics bit-identical; a 100 k-ray sample is bit-exact against the CPU oracle (its own tree over the same triangles);
  * linearity of the film: the sums of interleaved tile shares, and of sample slices, reproduce the full frame (tiles bit for bit);
  * determinism: the same frame rendered twice is bit-identical (the queues are compacted with atomics, the film is not);
  * every ray the stats count is a ray the wavefront could have traced: paths <= rays <= max_segments * paths.
"""
import numpy as np
import pytest

import checker
import common
from vermilion_b200 import engine, scenes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2():
    pos, nrm, uv = scenes.spheres(4, 3, 6, 150.0)
    assert len(pos) == 983040
    e = engine.Engine(device=0, max_paths=8 << 20)
    e.set_scene(pos, nrm, uv)
    yield e, pos, nrm, uv
    e.close()


def test_c2_two_traversals_and_oracle_sample(c2):
    e, pos, nrm, uv = c2
    box = (np.array([-1900, 10, -1900], np.float32), np.array([1900, 990, 1900], np.float32))
    rays = np.concatenate([common.aimed_rays(pos, 1_000_000, 3, origin_box=box), common.aimed_rays(pos, 1_000_000, 4, spread=0.05, origin_box=box)])
    a = e.intersect(rays)
    b = e.intersect(rays, engine.ISECT_BVH2)
    same = a["tri"] == b["tri"]
    assert (~same).sum() <= 2e-5 * len(rays)
    assert a[same].tobytes() == b[same].tobytes()
    assert (a["tri"] != engine.MISS).mean() > 0.5
    sub = np.concatenate([rays[:50_000], rays[-50_000:]])
    ref = checker.Oracle(pos, nrm, uv).intersect(sub)
    c = common.classify_hits(np.concatenate([a[:50_000], a[-50_000:]]), ref)
    assert c["bad"] == 0 and c["tie"] + c["edge"] <= 5, c
    assert c["t_exact"] == c["hits"] and c["uv_err_max"] == 0.0, c


def test_c2_film_linearity_and_determinism(c2):
    e, pos, nrm, uv = c2
    cam = engine.camera_from_settings(1920, 1080, (0.0, 500.0, 1900.0), (0.0, 0.0, 0.0), 6.0, (3.6, 3.6 * 1080 / 1920), 4)
    e.set_camera(cam)
    e.render(mode=engine.MODE_FIXED, seed=1)
    full = e.read_image(engine.FILM_SUM)
    st = e.stats()
    npx = 1920 * 1080
    assert st["paths_traced"] == 4 * npx and 4 * npx <= st["rays_traced"] <= 5 * 4 * npx
    assert (full[:, 3] == 4).all() and np.isfinite(full).all() and full[:, :3].max() > 0
    e.render(mode=engine.MODE_FIXED, seed=1)
    assert np.array_equal(e.read_image(engine.FILM_SUM), full)
    acc = np.zeros_like(full)
    for r in range(4):  # the 4-GPU tile split of C4, on one device
        e.render(mode=engine.MODE_FIXED, seed=1, tile_size=64, tile_rank=r, tile_world=4)
        part = e.read_image(engine.FILM_SUM)
        assert ((part[:, 3] == 0) | (part == full).all(1)).all()
        acc += part
    assert np.array_equal(acc, full)
    acc = np.zeros_like(full)
    for s0, s1 in ((0, 1), (1, 4)):  # sample slices (the multi-GPU split of bench.py)
        e.render(mode=engine.MODE_FIXED, seed=1, sample_begin=s0, sample_end=s1)
        acc += e.read_image(engine.FILM_SUM)
    assert np.array_equal(acc[:, 3], full[:, 3]) and np.allclose(acc, full, rtol=1e-6, atol=1e-6)
    e.render(mode=engine.MODE_FIXED, seed=2)
    assert not np.array_equal(e.read_image(engine.FILM_SUM), full)  # the seed is part of the stream


def test_c2_bruteforce_first_hit_depth_matches_intersect(c2):
    """BRUTEFORCE at full size: the depth plane (distance of the last primary RayCast, integrators.cpp:181) is consistent with the film:
    every pixel took 3 samples (the abs(int) early-out), alpha = 1 inside the closed room, depth finite and positive."""
    e, pos, nrm, uv = c2
    cam = engine.camera_from_settings(1920, 1080, (-900.0, 700.0, 1800.0), (-10.0, 25.0, 0.0), 6.0, (3.6, 3.6 * 1080 / 1920), 128)
    e.set_camera(cam)
    e.render(mode=engine.MODE_BRUTEFORCE, seed=1)
    img = e.read_image()
    film = e.read_image(engine.FILM_SUM)
    assert (film[:, 3] == 3).all()
    assert (img[:, 3] == 1).all() and np.isfinite(img[:, 4]).all() and (img[:, 4] > 0).all()
    assert (img[:, :3] >= 0).all() and (img[:, :3] <= 1).all()
    e.render(mode=engine.MODE_BRUTEFORCE, seed=1)
    assert np.array_equal(e.read_image(), img)

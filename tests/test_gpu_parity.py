"""GPU parity tests (-m gpu): the CUDA path through the C ABI against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star / SURVEY.md §8c):
  * closest-hit triangle id: exact, except equal-t ties and border cases, which are counted and bounded;
  * t, u, v where ids agree: bit-exact (the triangle test restates the reference arithmetic without FMA);
  * RayCast records: bit-exact wherever the id agrees;
  * camera rays: bit-exact;
  * Radiance with the shared Philox stream: >= 99 % of paths bit-exact (libm cos/sin may differ in the last ulp);
  * images: RMSE thresholds written in each test.
"""
import numpy as np
import pytest

import checker
import common
from vermilion_b200 import engine, scenes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    e = engine.Engine(device=0)
    yield e
    e.close()


def scene_soup(n=100_000):
    return scenes.soup(n, seed=1234)


def scene_spheres():
    return scenes.spheres(2, 2, 4, 150.0)  # 4 x 5120


def scene_cornell():
    return scenes.cornell()


SCENES = {"soup": scene_soup, "spheres": scene_spheres, "cornell": scene_cornell}


def rays_for(name, pos, n, seed):
    if name == "soup":
        a = scenes.uniform_rays(n // 2, seed=seed, lo=0.0, hi=1000.0).view(checker.RAY).reshape(-1)
        b = common.aimed_rays(pos, n - n // 2, seed + 1)
        return np.concatenate([a, b])
    box = (np.array([-1900, 10, -1900], np.float32), np.array([1900, 990, 1900], np.float32))
    a = common.aimed_rays(pos, n // 2, seed, origin_box=box)
    b = common.aimed_rays(pos, n - n // 2, seed + 1, spread=0.05, origin_box=box)
    return np.concatenate([a, b])


@pytest.mark.parametrize("name", ["soup", "spheres", "cornell"])
@pytest.mark.parametrize("flags", [0, engine.ISECT_BVH2])
def test_intersect_matches_oracle(eng, name, flags):
    pos, nrm, uv = SCENES[name]()
    rays = rays_for(name, pos, 200_000, 11)
    eng.set_scene(pos, nrm, uv)
    got = eng.intersect(rays, flags)
    ref = checker.Oracle(pos, nrm, uv).intersect(rays)
    c = common.classify_hits(got, ref)
    print(name, flags, c)
    assert c["bad"] == 0, c
    assert c["tie"] + c["edge"] <= max(20, 2e-4 * c["n"]), c  # reported, bounded
    assert c["hits"] > 0.3 * c["n"]
    assert c["t_exact"] == c["hits"], c        # bit-exact t
    assert c["uv_err_max"] == 0.0, c           # bit-exact barycentrics
    miss = got["tri"] == engine.MISS
    assert np.all(got["t"][miss] == np.float32(999999999.0)) and np.all(got["u"][miss] == 0)


def test_raycast_matches_oracle(eng):
    pos, nrm, uv = scene_spheres()
    rays = rays_for("spheres", pos, 100_000, 5)
    eng.set_scene(pos, nrm, uv)
    got = eng.raycast(rays)
    ref = checker.Oracle(pos, nrm, uv).raycast(rays)
    same = got["tri"] == ref["tri"]
    assert same.mean() > 0.9998
    for f in ("P", "dist", "N", "flags", "uv", "Le"):
        a, b = got[f][same], ref[f][same]
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), f
    assert (ref["flags"] & 1).all()  # inside the room every ray hits something


def test_raycast_without_triangles_on_ray(eng):
    """Sphere-only hits (no material) and NaN rays (the reference's dead diffuse bounces)."""
    pos, nrm, uv = scene_cornell()
    o = np.tile(np.array([[1000.0, 500.0, 1000.0]], np.float32), (1000, 1))
    d = np.random.default_rng(3).normal(size=(1000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[::50] = np.nan
    rays = checker.make_rays(o, d)
    eng.set_scene(pos, nrm, uv)
    got = eng.raycast(rays)
    ref = checker.Oracle(pos, nrm, uv).raycast(rays)
    assert np.array_equal(got["flags"], ref["flags"])
    ok = ~np.isnan(d[:, 0])
    for f in ("P", "dist", "N", "uv", "Le"):
        assert np.array_equal(got[f][ok].view(np.uint32), ref[f][ok].view(np.uint32)), f
    assert (got["flags"][~ok] == 0).all()


def test_occluded_matches_oracle(eng):
    pos, nrm, uv = scene_spheres()
    rays = rays_for("spheres", pos, 50_000, 9)
    eng.set_scene(pos, nrm, uv)
    got = eng.occluded(rays)
    ref = checker.Oracle(pos, nrm, uv).occluded(rays)
    assert (got != ref).sum() <= 10
    assert 0.2 < ref.mean() <= 1.0


def test_camera_rays_bit_exact(eng):
    cam = engine.camera_from_settings(648, 432, (-4000, 1600, 8000), (0, 25, 0), 6.0, (3.6, 2.4), 128)  # renderEngine.cpp:119-142
    eng.set_camera(cam)
    rng = np.random.default_rng(1)
    n = 20000
    px = rng.integers(0, 648 * 432, n).astype(np.uint64)
    sx = rng.integers(0, 2, n).astype(np.uint32)
    sy = rng.integers(0, 2, n).astype(np.uint32)
    jx = (rng.random(n) * 0.5).astype(np.float32)
    jy = (rng.random(n) * 0.5).astype(np.float32)
    got = eng.camera_rays(px, sx, sy, jx, jy)
    ocam = checker.OrcCamera(cam.width, cam.height, cam.position, cam.rotation, cam.dist_to_film, cam.sensor_x, cam.sensor_y, cam.spp)
    for i in range(0, n, 7):
        ref = checker.Oracle.camera_ray(ocam, int(px[i]), int(sx[i]), int(sy[i]), float(jx[i]), float(jy[i]))
        assert got[i:i + 1].tobytes() == ref.tobytes(), i


@pytest.mark.parametrize("mode", [checker.MODE_COMPAT, checker.MODE_FIXED])
def test_radiance_matches_oracle(eng, mode):
    pos, nrm, uv = scene_spheres()
    rays = rays_for("spheres", pos, 40_000, 21)
    eng.set_scene(pos, nrm, uv)
    got, cnt = eng.radiance(rays, seed=77, mode=mode)
    ref, rcnt = checker.Oracle(pos, nrm, uv).radiance(rays, 77, checker.RNG_PHILOX, mode)
    exact = (got.view(np.uint32) == ref.view(np.uint32)).all(1) & (cnt == rcnt)
    print("radiance exact fraction", exact.mean(), "mean rays/path", cnt.mean(), rcnt.mean())
    assert exact.mean() > 0.99
    assert abs(cnt.mean() - rcnt.mean()) < 0.01 * rcnt.mean()
    assert abs(got[:, :3].mean() - ref[:, :3].mean()) < 0.02 * max(ref[:, :3].mean(), 1e-3)


def _ocam(cam):
    return checker.OrcCamera(cam.width, cam.height, cam.position, cam.rotation, cam.dist_to_film, cam.sensor_x, cam.sensor_y, cam.spp)


def test_render_fixed_matches_oracle(eng):
    pos, nrm, uv = scene_spheres()
    cam = engine.camera_from_settings(96, 64, (0, 500, 1900), (0, 0, 0), 6.0, (3.6, 2.4), 16)
    eng.set_scene(pos, nrm, uv)
    eng.set_camera(cam)
    eng.render(mode=engine.MODE_FIXED, seed=9)
    img = eng.read_image()
    st = eng.stats()
    ref, nrays = checker.Oracle(pos, nrm, uv).render(_ocam(cam), checker.MODE_FIXED, checker.RNG_PHILOX, 9)
    assert np.array_equal(img[:, 3:], ref[:, 3:])  # alpha and sample counts
    exact = (img[:, :3].view(np.uint32) == ref[:, :3].view(np.uint32)).all(1).mean()
    rmse = float(np.sqrt(np.mean((img[:, :3] - ref[:, :3]) ** 2)))
    print("fixed: exact pixels", exact, "rmse", rmse, "rays", st["rays_traced"], nrays)
    assert exact > 0.97
    assert rmse < 5e-3
    assert abs(st["rays_traced"] - nrays) <= 0.002 * nrays


def test_render_compat_matches_oracle(eng):
    pos, nrm, uv = scene_cornell()
    cam = engine.camera_from_settings(64, 64, (0, 275, 1100), (0, 0, 0), 6.0, (3.6, 3.6), 16)
    eng.set_scene(pos, nrm, uv)
    eng.set_camera(cam)
    eng.render(mode=engine.MODE_COMPAT, seed=4)
    img = eng.read_image()
    ref, nrays = checker.Oracle(pos, nrm, uv).render(_ocam(cam), checker.MODE_COMPAT, checker.RNG_PHILOX, 4)
    same_n = (img[:, 4] == ref[:, 4]).mean()
    rmse = float(np.sqrt(np.mean((img[:, :3] - ref[:, :3]) ** 2)))
    print("compat: same sample count", same_n, "rmse", rmse)
    assert same_n > 0.97
    assert rmse < 1e-2
    assert abs(eng.stats()["rays_traced"] - nrays) <= 0.01 * nrays


def test_tile_and_sample_split_sum_to_full_frame(eng):
    pos, nrm, uv = scene_spheres()
    cam = engine.camera_from_settings(80, 48, (0, 500, 1900), (0, 0, 0), 6.0, (3.6, 2.16), 8)
    eng.set_scene(pos, nrm, uv)
    eng.set_camera(cam)
    eng.render(mode=engine.MODE_FIXED, seed=2)
    full = eng.read_image(engine.FILM_SUM)
    acc = np.zeros_like(full)
    for r in range(3):  # interleaved 16x16 tiles over 3 ranks
        eng.render(mode=engine.MODE_FIXED, seed=2, tile_size=16, tile_rank=r, tile_world=3)
        acc += eng.read_image(engine.FILM_SUM)
    assert np.array_equal(acc, full)
    acc = np.zeros_like(full)
    for s0, s1 in ((0, 3), (3, 8)):  # sample slices
        eng.render(mode=engine.MODE_FIXED, seed=2, sample_begin=s0, sample_end=s1)
        acc += eng.read_image(engine.FILM_SUM)
    assert np.allclose(acc, full, rtol=1e-6, atol=1e-6) and np.array_equal(acc[:, 3], full[:, 3])


def test_edge_cases(eng):
    # empty scene: every ray misses; spheres still answer RayCast
    eng.set_scene(np.zeros((0, 9), np.float32))
    rays = scenes.uniform_rays(1000, seed=1, lo=-100.0, hi=900.0)
    h = eng.intersect(rays)
    assert (h["tri"] == engine.MISS).all() and (h["t"] == np.float32(999999999.0)).all()
    assert len(eng.intersect(rays[:0])) == 0
    # one triangle, ragged ray counts, axis-parallel and zero-component directions
    tri = np.array([[0, 0, 0, 10, 0, 0, 0, 10, 0]], np.float32)
    eng.set_scene(tri, spheres=np.zeros(0, engine.SPHERE))
    orc = checker.Oracle(tri, spheres=False)
    o = np.array([[2, 2, 5], [2, 2, -5], [20, 2, 5], [2, 2, 5], [0, 0, 5], [5, 5, 5]], np.float32)
    d = np.array([[0, 0, -1], [0, 0, 1], [0, 0, -1], [0, 0, 1], [0, 0, -1], [0, 0, -1]], np.float32)
    for n in (1, 5, 6):
        r = checker.make_rays(o[:n], d[:n])
        assert eng.intersect(r).tobytes() == orc.intersect(r).tobytes()
    # degenerate (zero-area) triangles never hit and do not break the build
    deg = np.zeros((40, 9), np.float32)
    deg[:, :] = np.arange(40, dtype=np.float32)[:, None]
    eng.set_scene(np.concatenate([deg, tri]), spheres=np.zeros(0, engine.SPHERE))
    r = checker.make_rays(o, d)
    got = eng.intersect(r)
    ref = checker.Oracle(np.concatenate([deg, tri]), spheres=False).intersect(r)
    # ray 4 hits exactly the vertex (0,0,0) with d.x = d.y = 0: the reference's slab test computes 0*inf = NaN there and drops
    # the hit (bbox.cpp:70-83); that is a border case (u = v = 0), reported by classify_hits, everything else is bit-exact
    c = common.classify_hits(got, ref)
    assert c["bad"] == 0 and c["tie"] + c["edge"] <= 1, c
    keep = np.array([0, 1, 2, 3, 5])
    assert got[keep].tobytes() == ref[keep].tobytes()


def test_product_path_fails_loudly_without_commit(eng):
    e2 = engine.Engine(device=0)
    with pytest.raises(engine.VmnError):
        e2.intersect(np.zeros(1, checker.RAY))
    e2.close()


def test_rgba8_pack_matches_saveframe_conversion(eng):
    """vmn_read_image_rgba8 against the conversion loop of Camera::saveFrame (camera.cpp:138-175) applied to the RGBAZ film."""
    pos, nrm, uv = scene_spheres()
    cam = engine.camera_from_settings(96, 64, (0, 500, 1900), (0, 0, 0), 6.0, (3.6, 2.4), 16)
    eng.set_scene(pos, nrm, uv)
    eng.set_camera(cam)
    eng.render(mode=engine.MODE_FIXED, seed=9, tile_size=16, tile_rank=0, tile_world=2)  # half the tiles stay unrendered (all zero)
    img = eng.read_image()
    rgba, depth = eng.read_image_rgba8()
    want = np.floor(img[:, :4] * np.float32(255)).astype(np.uint8)
    assert np.array_equal(rgba, want)
    assert np.array_equal(depth, img[:, 4])
    assert rgba[:, 3].max() == 255 and rgba[:, 3].min() == 0 and rgba[:, :3].max() > 0


def test_progressive_accumulate_equals_one_shot(eng):
    """VMN_RENDER_ACCUMULATE: sample slices added into the film reproduce the one-shot sums (same per-pixel sample order)."""
    pos, nrm, uv = scene_spheres()
    cam = engine.camera_from_settings(64, 48, (0, 500, 1900), (0, 0, 0), 6.0, (3.6, 2.7), 12)
    eng.set_scene(pos, nrm, uv)
    eng.set_camera(cam)
    eng.render(mode=engine.MODE_FIXED, seed=5)
    full = eng.read_image(engine.FILM_SUM)
    for i, (a, b) in enumerate(((0, 5), (5, 6), (6, 12))):
        eng.render(mode=engine.MODE_FIXED, seed=5, sample_begin=a, sample_end=b, accumulate=i > 0)
    assert np.array_equal(eng.read_image(engine.FILM_SUM), full)
    with pytest.raises(engine.VmnError):
        eng.render(mode=engine.MODE_COMPAT, seed=5, accumulate=True)


@pytest.mark.parametrize("channels", [1, 3, 4])
def test_textured_radiance_and_render_match_oracle(channels):
    """N4: MeshEngine::boundTextures[0] scales the throughput at every diffuse triangle bounce (pathtracer.cpp:62-80,
    VermiTexture::Sample meshEngine.cpp:21-46). Also the path whose throughput buffer is written (kThrWritten in k_shade)."""
    pos, nrm, uv = scene_cornell()
    rng = np.random.default_rng(5)
    tex = rng.random((16, 8, channels), dtype=np.float32) * np.float32(1.5)
    tex[3, 2] = np.float32(np.inf)  # a non-finite texel: inf * 0 = NaN must reach the radiance exactly where the reference's does
    rays = rays_for("cornell", pos, 60_000, 23)
    orc = checker.Oracle(pos, nrm, uv)
    orc.set_texture(tex)
    cam = engine.camera_from_settings(64, 64, (0, 275, 1100), (0, 0, 0), 6.0, (3.6, 3.6), 16)
    with engine.Engine(device=0) as e:
        e.set_scene(pos, nrm, uv, texture=tex)
        for mode in (checker.MODE_COMPAT, checker.MODE_FIXED):
            got, cnt = e.radiance(rays, 31, mode)
            want, wcnt = orc.radiance(rays, 31, checker.RNG_PHILOX, mode)
            same = ((got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))).all(1)
            print("textured radiance mode", mode, "bit-exact", same.mean(), "nan paths", np.isnan(want).any(1).mean())
            assert same.mean() >= 0.99 and (cnt == wcnt).mean() >= 0.99
        e.set_camera(cam)
        e.render(mode=engine.MODE_FIXED, seed=9)
        img = e.read_image()
    ref, _ = orc.render(_ocam(cam), checker.MODE_FIXED, checker.RNG_PHILOX, 9)
    same = ((img.view(np.uint32) == ref.view(np.uint32)) | (np.isnan(img) & np.isnan(ref))).all(1)
    print("textured render bit-exact pixels", same.mean())
    assert same.mean() > 0.97
    plain, _ = checker.Oracle(pos, nrm, uv).render(_ocam(cam), checker.MODE_FIXED, checker.RNG_PHILOX, 9)
    assert not np.array_equal(plain, ref)  # the texture changes the picture


def test_greedy_collapse_knob_keeps_hits(monkeypatch):
    """VMN_COLLAPSE=greedy (the round-1 collapse rule, kept as an A/B knob of the builder): another tree, the same closest hits."""
    pos, nrm, uv = scenes.spheres(2, 2, 4, 150.0)
    box = (np.array([-1900, 10, -1900], np.float32), np.array([1900, 990, 1900], np.float32))
    rays = common.aimed_rays(pos, 100_000, 21, spread=0.05, origin_box=box)
    with engine.Engine(device=0, counters=True) as e:
        e.set_scene(pos, nrm, uv)
        a = e.intersect(rays)
        dp_stats = e.stats()
        monkeypatch.setenv("VMN_COLLAPSE", "greedy")
        e.set_scene(pos, nrm, uv)
        b = e.intersect(rays)
        greedy_stats = e.stats()
    same = a["tri"] == b["tri"]
    assert (~same).sum() <= 5 and a[same].tobytes() == b[same].tobytes()
    assert dp_stats["n_wide_nodes"] != greedy_stats["n_wide_nodes"]
    # the cost-table collapse must not test more triangles than the greedy one on this scene
    assert dp_stats["tris_tested"] <= greedy_stats["tris_tested"]


def test_read_bandwidth_probe_separates_l2_from_hbm():
    """vmn_probe_read_bandwidth (the L2 ceiling bench.py reports beside the HBM roofline): a 32 MiB working set is served by the
    L2 and must read faster than a 2 GiB one, which must itself reach a sane fraction of HBM bandwidth; bad arguments are refused."""
    with engine.Engine(device=0) as eng:
        l2 = eng.probe_read_bandwidth(32 << 20, 100)
        hbm = eng.probe_read_bandwidth(2 << 30, 2)
        assert l2 > hbm > 2000.0, (l2, hbm)
        with pytest.raises(Exception):
            eng.probe_read_bandwidth(1024, 1)


@pytest.mark.parametrize("flags", [0, engine.ISECT_BVH2])
def test_signed_zero_direction_components_match_oracle(eng, flags):
    """Axis-parallel rays whose zero components carry either sign. The reference's inv_d is +-inf there (Ray.h:8) and its slab test
    passes for an origin inside the slab whatever the sign of the zero (bbox.cpp:70-83); the wide traversal picks the near plane from
    the direction's sign bit, which must agree with the sign of the clamped reciprocal it multiplies with."""
    pos, nrm, uv = scene_spheres()
    eng.set_scene(pos, nrm, uv)
    rng = np.random.default_rng(17)
    n = 60_000
    tri = np.asarray(pos, np.float32).reshape(-1, 3, 3)
    tgt = tri[rng.integers(0, len(tri), n)].mean(1)
    axis = rng.integers(0, 3, n)
    sign = rng.choice(np.array([-1.0, 1.0], np.float32), n)
    d = np.zeros((n, 3), np.float32)
    d[np.arange(n), axis] = sign
    zero_sign = rng.choice(np.array([-0.0, 0.0], np.float32), (n, 3))
    d = np.where(d == 0, zero_sign, d).astype(np.float32)
    o = tgt.copy()
    o[np.arange(n), axis] -= sign * np.float32(700.0)
    rays = checker.make_rays(o.astype(np.float32), d)
    assert np.signbit(rays["d"]).any() and (rays["d"] == 0).any()
    got = eng.intersect(rays, flags)
    ref = checker.Oracle(pos, nrm, uv).intersect(rays)
    c = common.classify_hits(got, ref)
    print(flags, c)
    assert c["bad"] == 0, c
    assert c["tie"] + c["edge"] <= 0.01 * n, c  # rays through triangle centroids of a closed mesh rarely graze an edge
    assert c["hits"] > 0.9 * n and c["t_exact"] == c["hits"], c


@pytest.mark.parametrize("textured", [False, True])
def test_one_launch_frames_equal_the_wavefront_bit_for_bit(monkeypatch, textured):
    """COMPAT and BRUTEFORCE frames of up to 4 Mi pixels run as ONE kernel (k_mega_compat / k_mega_bf: the pixel loop, the early-out and
    the path loop in registers, no host round trips); larger ones, or VMN_MEGA=0, go through the wavefront stages. Both share the
    arithmetic (raycast_finish, radiance_bounce, bf_*), so films, sample counts and ray counts must be identical — with tiles, shadow
    rays and textures too."""
    pos, nrm, uv = scenes.spheres(2, 2, 3, 150.0)
    rng = np.random.default_rng(5)
    tex = (rng.random((8, 8, 3), dtype=np.float32) - np.float32(0.5)) * np.float32(0.6) if textured else None
    tex1 = rng.random((16, 8, 4), dtype=np.float32) if textured else None
    cam = engine.camera_from_settings(160, 120, (0.0, 500.0, 1900.0), (0.0, 5.0, 0.0), 6.0, (3.6, 2.7), 16)
    out = {}
    with engine.Engine(device=0, max_paths=1 << 16) as e:  # a small pool: the wavefront path runs in several chunks
        e.set_scene(pos, nrm, uv, texture=tex, texture1=tex1)
        e.set_camera(cam)
        for mega in ("1", "0"):
            monkeypatch.setenv("VMN_MEGA", mega)
            for name, kw in (("compat", dict(mode=engine.MODE_COMPAT)), ("compat_tiles", dict(mode=engine.MODE_COMPAT, tile_size=32, tile_rank=1, tile_world=3)),
                             ("bf", dict(mode=engine.MODE_BRUTEFORCE)), ("bf_shadow", dict(mode=engine.MODE_BRUTEFORCE, shadow=True)),
                             ("bf_tiles", dict(mode=engine.MODE_BRUTEFORCE, tile_size=32, tile_rank=2, tile_world=3))):
                e.render(seed=7, **kw)
                st = e.stats()
                out[(name, mega)] = (e.read_image().copy(), e.read_image(engine.FILM_SUM).copy(), st["rays_traced"], st["paths_traced"], st["kernel_launches"])
    for name in ("compat", "compat_tiles", "bf", "bf_shadow", "bf_tiles"):
        a, b = out[(name, "1")], out[(name, "0")]
        assert np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32)), name
        assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), name
        assert a[2] == b[2] and a[3] == b[3], (name, a[2:], b[2:])
        assert a[4] == 1 and b[4] > 10

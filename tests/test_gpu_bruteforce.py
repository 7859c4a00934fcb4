"""GPU parity tests (-m gpu) of VMN_MODE_BRUTEFORCE — BruteForceTracer::Render (reference core/integrators/integrators.cpp:9-186),
the RenderEngine's default integrator — through the C ABI against the CPU oracle drawing from the same Philox stream.

Bars: sample counts (film n) exact; every output float (clamped rgb, alpha = hits / samples, depth = last primary distance) bit-exact
for >= 99.5 % of pixels (the remainder are equal-t ties at box edges, where the GPU tree visits triangles in another order), NaN
pixels (zero normals) at the same places. The oracle itself is pinned bit-exactly to the real BruteForceTracer::Render
(tests/test_oracle_vs_ref.py, tests/golden/bruteforce.npz).
"""
import numpy as np
import pytest

import checker
from vermilion_b200 import engine, scenes

pytestmark = pytest.mark.gpu

CORNELL_CAM = dict(position=(0, 275, 1100), rotation_deg=(0, 0, 0), back_distance=6.0)


def _ocam(cam):
    return checker.OrcCamera(cam.width, cam.height, cam.position, cam.rotation, cam.dist_to_film, cam.sensor_x, cam.sensor_y, cam.spp)


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def bf_textures(kind, seed=3):
    rng = np.random.default_rng(seed)
    nmap = (rng.random((8, 8, 3), dtype=np.float32) - np.float32(0.5)) * np.float32(0.6)
    alb = rng.random((16, 8, 4), dtype=np.float32) * np.float32(60.0 if kind == "hot" else 1.0)
    return nmap, alb


def _compare(img, ref, min_exact=0.995):
    same = ((_bits(img) == _bits(ref)) | (np.isnan(img) & np.isnan(ref))).all(1)  # NaN payloads differ between x86 and the GPU
    print("bit-exact pixels", same.mean())
    assert same.mean() >= min_exact
    assert np.array_equal(np.isnan(img), np.isnan(ref))


@pytest.mark.parametrize("scene,textures,shadow,spheres,zero_normals", [
    ("cornell", None, False, True, False),
    ("cornell", "plain", False, True, False),
    ("cornell", "hot", False, True, False),
    ("cornell", "hot", True, True, False),
    ("cornell", None, True, True, False),
    ("grid", None, False, False, False),      # no room spheres: primary misses (alpha < 1, depth inf) and probe misses (vNDL*0.9+0.1)
    ("grid", "plain", True, False, False),
    ("cornell", None, False, True, True),     # zero normals: NaN colours through abs(int) and std::min/max
    ("city", "plain", True, True, False),
])
def test_bruteforce_matches_oracle(scene, textures, shadow, spheres, zero_normals):
    pos, nrm, uv = {"cornell": scenes.cornell, "city": lambda: scenes.city(3, 3, 4, 2024), "grid": lambda: scenes.spheres(4, 3, 3, 150.0)}[scene]()
    if zero_normals:
        nrm = None
    cam = {"cornell": engine.camera_from_settings(96, 80, (0, 275, 1100), (0, 0, 0), 6.0, (3.6, 3.0), 12),
           "city": engine.camera_from_settings(96, 80, (-1500, 900, 1800), (-20, 30, 0), 6.0, (3.6, 3.0), 12),
           "grid": engine.camera_from_settings(96, 80, (0, 500, 1900), (0, 0, 0), 6.0, (3.6, 3.0), 12)}[scene]
    orc = checker.Oracle(pos, nrm, uv, spheres=spheres)
    t0 = t1 = None
    if textures:
        t0, t1 = bf_textures(textures)
        orc.set_texture(t0)
        orc.add_texture(t1)
    with engine.Engine(device=0) as eng:
        eng.set_scene(pos, nrm, uv, spheres=None if spheres else np.zeros(0, engine.SPHERE), texture=t0, texture1=t1)
        eng.set_camera(cam)
        eng.render(mode=engine.MODE_BRUTEFORCE, seed=11, shadow=shadow)
        img = eng.read_image()
        film = eng.read_image(engine.FILM_SUM)
        st = eng.stats()
    ref, taken, nq = orc.render_bruteforce(_ocam(cam), rng=checker.RNG_PHILOX, seed=11, flags=checker.BF_SHADOW if shadow else 0)
    assert (film[:, 3] == taken).mean() >= 0.995
    _compare(img, ref)
    assert abs(int(st["rays_traced"]) - int(nq)) <= 0.01 * nq
    assert st["paths_traced"] >= taken.sum() * 0.99
    if textures == "hot":
        assert taken.max() > 3
    if not spheres:
        assert np.isinf(ref[:, 4]).any() and (ref[:, 3] < 1).any()
    if zero_normals:
        assert np.isnan(ref[:, :3]).any()
    if shadow and scene == "cornell":
        plain, _, _ = orc.render_bruteforce(_ocam(cam), rng=checker.RNG_PHILOX, seed=11, flags=0)
        assert not np.array_equal(_bits(plain), _bits(ref))  # the shadow ray changes the picture


def test_bruteforce_tiles_sum_to_full_frame_and_pack():
    pos, nrm, uv = scenes.cornell()
    cam = engine.camera_from_settings(80, 48, (0, 275, 1100), (0, 0, 0), 6.0, (3.6, 2.16), 8)
    with engine.Engine(device=0, max_paths=1024) as eng:  # 1024 paths in flight: the frame is rendered in four chunks
        eng.set_scene(pos, nrm, uv)
        eng.set_camera(cam)
        eng.render(mode=engine.MODE_BRUTEFORCE, seed=2)
        full = eng.read_image()
        rgba, depth = eng.read_image_rgba8()
        want = np.floor(full[:, :4] * np.float32(255)).astype(np.uint8)  # Camera::saveFrame (camera.cpp:138-175)
        assert np.array_equal(rgba, want) and np.array_equal(depth, full[:, 4])
        acc = np.zeros_like(full)
        for r in range(3):
            eng.render(mode=engine.MODE_BRUTEFORCE, seed=2, tile_size=16, tile_rank=r, tile_world=3)
            part = eng.read_image()
            assert ((part == 0).all(1) | (part == full).all(1)).all()  # a pixel is either not mine or identical
            acc += part
        assert np.array_equal(acc, full)
    ref, _, _ = checker.Oracle(pos, nrm, uv).render_bruteforce(_ocam(cam), rng=checker.RNG_PHILOX, seed=2)
    _compare(full, ref)


def test_bruteforce_argument_errors():
    pos, nrm, uv = scenes.cornell()
    with engine.Engine(device=0) as eng:
        eng.set_scene(pos, nrm, uv)
        eng.set_camera(engine.camera_from_settings(8, 8, (0, 275, 1100), (0, 0, 0), 6.0, (3.6, 3.6), 70000))
        with pytest.raises(engine.VmnError):  # the reference's uint16_t sample loop never terminates above 65535 spp
            eng.render(mode=engine.MODE_BRUTEFORCE)
        eng.set_camera(engine.camera_from_settings(8, 8, (0, 275, 1100), (0, 0, 0), 6.0, (3.6, 3.6), 4))
        with pytest.raises(engine.VmnError):
            eng.render(mode=engine.MODE_BRUTEFORCE, accumulate=True)
        with pytest.raises(engine.VmnError):
            eng.set_texture(2, np.zeros((2, 2, 3), np.float32))
        eng.render(mode=engine.MODE_BRUTEFORCE)
        assert eng.read_image().shape == (64, 5)


def test_empty_scene_renders_only_the_room_spheres():
    """No triangles at all: MeshEngine::RayCast still answers with the eight spheres; both integrators match the oracle bit for bit."""
    cam = engine.camera_from_settings(48, 32, (0, 500, 1900), (0, 0, 0), 6.0, (3.6, 2.4), 8)
    orc = checker.Oracle(np.zeros((0, 9), np.float32))
    with engine.Engine(device=0) as eng:
        eng.set_scene(np.zeros((0, 9), np.float32))
        eng.set_camera(cam)
        eng.render(mode=engine.MODE_BRUTEFORCE, seed=3)
        bf = eng.read_image()
        eng.render(mode=engine.MODE_FIXED, seed=3)
        pt = eng.read_image()
        eng.render(mode=engine.MODE_COMPAT, seed=3)
        pc = eng.read_image()
    ref, _, _ = orc.render_bruteforce(_ocam(cam), seed=3)
    assert np.array_equal(_bits(bf), _bits(ref))
    want, _ = orc.render(_ocam(cam), checker.MODE_FIXED, checker.RNG_PHILOX, 3)
    assert np.array_equal(_bits(pt), _bits(want))
    want, _ = orc.render(_ocam(cam), checker.MODE_COMPAT, checker.RNG_PHILOX, 3)
    assert np.array_equal(_bits(pc), _bits(want))

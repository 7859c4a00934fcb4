"""The drop-in seam: Vermilion::B200PathTracer (vermilion_b200/host/b200_pathtracer.cpp) called through the reference's own
Integrator vtable with a reference Camera and a reference MeshEngine (oracle/_ref/libvermadapter.so, built from the reference
headers/objects where /root/reference exists; the prebuilt .so travels to the GPU box)."""
import ctypes as C
import os

import numpy as np
import pytest

import checker
from vermilion_b200 import scenes

ADAPTER = os.path.join(checker.ORACLE_DIR, "_ref", "libvermadapter.so")
needs_adapter = pytest.mark.skipif(not (os.path.exists(ADAPTER) and checker.Ref.available("ieee")), reason="oracle/_ref/libvermadapter.so not built")
RGB, RGBA, RGBAZ, DEPTH = 0, 1, 2, 3  # vermRenderMode (camera.h:21-30)


def adapter_render(ref, cs, render_mode, mode, seed, max_segments=5):
    L = C.CDLL(ADAPTER)
    L.adapter_render.argtypes = [C.c_void_p, C.POINTER(checker.RefCameraSettings), C.c_int, C.c_int, C.c_uint64, C.c_uint32, C.c_void_p,
                                 C.POINTER(C.c_uint64), C.c_char_p, C.c_int]
    per = {RGB: 3, RGBA: 4, RGBAZ: 5, DEPTH: 1}[render_mode]
    img = np.zeros((cs.width * cs.height, per), np.float32)
    rays = C.c_uint64()
    err = C.create_string_buffer(512)
    rc = L.adapter_render(ref.h, C.byref(cs), render_mode, mode, seed, max_segments, img.ctypes.data_as(C.c_void_p), C.byref(rays), err, 512)
    return rc, img, rays.value, err.value.decode()


@needs_adapter
def test_adapter_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    pos, nrm, uv = scenes.cornell()
    ref = checker.Ref(pos, nrm, uv)
    cs = checker.camera_settings(16, 16, (0, 275, 1100), (0, 0, 0), 6.0, (3.6, 3.6), 16)
    rc, img, _, err = adapter_render(ref, cs, RGBAZ, checker.MODE_FIXED, 3)
    assert rc != 0 and "CUDA" in err
    assert not img.any()  # film untouched: nothing was rendered on the CPU instead


@pytest.mark.gpu
@needs_adapter
@pytest.mark.parametrize("mode", [checker.MODE_COMPAT, checker.MODE_FIXED])
def test_adapter_matches_oracle_through_the_integrator_interface(mode):
    from vermilion_b200 import engine
    pos, nrm, uv = scenes.spheres(2, 2, 3, 150.0)
    ref = checker.Ref(pos, nrm, uv)
    cs = checker.camera_settings(64, 48, (0, 500, 1900), (0, 5, 0), 6.0, (3.6, 2.7), 16)
    rc, img, nrays, err = adapter_render(ref, cs, RGBAZ, mode, 11)
    assert rc == 0, err
    # the same frame through the plain C ABI: identical bits
    with engine.Engine(device=0) as eng:
        eng.set_scene(pos, nrm, uv)
        eng.set_camera(engine.camera_from_settings(64, 48, (0, 500, 1900), (0, 5, 0), 6.0, (3.6, 2.7), 16))
        eng.render(mode=mode, seed=11)
        direct = eng.read_image()
    assert np.array_equal(img.view(np.uint32), direct.view(np.uint32))
    # and against the CPU oracle with the shared Philox stream
    oimg, onrays = checker.Oracle(pos, nrm, uv).render(ref.camera_fields(cs), mode, checker.RNG_PHILOX, 11)
    assert np.array_equal(img[:, 3:], oimg[:, 3:])
    assert (img[:, :3].view(np.uint32) == oimg[:, :3].view(np.uint32)).all(1).mean() > 0.97
    assert abs(nrays - onrays) <= 0.002 * onrays
    # other film layouts go through Camera::setPixelValue (camera.cpp:88-124)
    rc, rgb, _, err = adapter_render(ref, cs, RGB, mode, 11)
    assert rc == 0 and np.array_equal(rgb, img[:, :3])
    rc, dep, _, err = adapter_render(ref, cs, DEPTH, mode, 11)  # the one-float-per-pixel film (camera.cpp:73-75, 114-116)
    assert rc == 0 and np.array_equal(dep[:, 0], img[:, 4])


@pytest.mark.gpu
@needs_adapter
@pytest.mark.parametrize("shadow", [False, True])
def test_bruteforce_adapter_matches_oracle_through_the_integrator_interface(shadow):
    """Vermilion::B200BruteForceTracer — the drop-in for the RenderEngine's default integrator (renderEngine.cpp:51) — with the
    reference's own MeshEngine holding two bound textures (normal map, albedo)."""
    pos, nrm, uv = scenes.cornell()
    ref = checker.Ref(pos, nrm, uv)
    orc = checker.Oracle(pos, nrm, uv)
    rng = np.random.default_rng(3)
    nmap = (rng.random((8, 8, 3), dtype=np.float32) - np.float32(0.5)) * np.float32(0.6)
    alb = rng.random((16, 8, 4), dtype=np.float32) * np.float32(60.0)
    ref.set_texture(nmap)
    ref.set_texture(alb)
    orc.set_texture(nmap)
    orc.add_texture(alb)
    cs = checker.camera_settings(64, 48, (0, 275, 1100), (0, 0, 0), 6.0, (3.6, 2.7), 12)
    rc, img, nrays, err = adapter_render(ref, cs, RGBAZ, 2 | (0x100 if shadow else 0), 11)
    assert rc == 0, err
    oimg, taken, nq = orc.render_bruteforce(ref.camera_fields(cs), rng=checker.RNG_PHILOX, seed=11, flags=checker.BF_SHADOW if shadow else 0)
    assert (img.view(np.uint32) == oimg.view(np.uint32)).all(1).mean() >= 0.995
    assert taken.max() > 3 and abs(nrays - nq) <= 0.01 * nq
    rc, rgba, _, err = adapter_render(ref, cs, RGBA, 2 | (0x100 if shadow else 0), 11)  # through Camera::setPixelValue
    assert rc == 0 and np.array_equal(rgba.view(np.uint32), img[:, :4].view(np.uint32))


@pytest.mark.gpu
@needs_adapter
def test_adapter_progressive_checkpoints_resume(tmp_path):
    """B200Options::checkpointPath: slices of 4 spp with a checkpoint after each; a run that finds the checkpoint of an interrupted
    render (here: one cut down to its first 8 of 16 spp) finishes it, and both equal the one-shot frame bit for bit."""
    from vermilion_b200 import film_io
    pos, nrm, uv = scenes.spheres(2, 2, 3, 150.0)
    ref = checker.Ref(pos, nrm, uv)
    cs = checker.camera_settings(64, 48, (0, 500, 1900), (0, 5, 0), 6.0, (3.6, 2.7), 16)
    rc, one_shot, _, err = adapter_render(ref, cs, RGBAZ, checker.MODE_FIXED, 11)
    assert rc == 0, err
    L = C.CDLL(ADAPTER)
    L.adapter_set_checkpoint.argtypes = [C.c_char_p, C.c_uint32]
    path = str(tmp_path / "film.vmn").encode()
    try:
        L.adapter_set_checkpoint(path, 4)
        rc, img, full_rays, err = adapter_render(ref, cs, RGBAZ, checker.MODE_FIXED, 11)
        assert rc == 0, err
        assert np.array_equal(img.view(np.uint32), one_shot.view(np.uint32))
        film, hdr = film_io.checkpoint_load(path.decode())
        assert hdr.frame_fingerprint != 0
        assert hdr.spp_done == 16 and hdr.spp_total == 16 and hdr.seed == 11
        # rebuild the state of a run interrupted after 8 spp and let the adapter resume it
        from vermilion_b200 import engine
        with engine.Engine(device=0) as eng:
            eng.set_scene(pos, nrm, uv)
            eng.set_camera(engine.camera_from_settings(64, 48, (0, 500, 1900), (0, 5, 0), 6.0, (3.6, 2.7), 16))
            eng.render(mode=engine.MODE_FIXED, seed=11, sample_begin=0, sample_end=8)
            half = eng.read_image(engine.FILM_SUM)
        film_io.checkpoint_save(path.decode(), half, 64, 48, 8, 16, 11, frame_fingerprint=hdr.frame_fingerprint)
        rc, img, nrays, err = adapter_render(ref, cs, RGBAZ, checker.MODE_FIXED, 11)
        assert rc == 0, err
        assert np.array_equal(img.view(np.uint32), one_shot.view(np.uint32))
        assert film_io.checkpoint_info(path.decode()).spp_done == 16
        assert nrays < 0.75 * full_rays  # only the second half was traced
        # a checkpoint of ANOTHER frame with the same size / spp / seed (here: a finished film full of garbage whose fingerprint
        # does not match this camera and scene) must be ignored, not returned as the image
        film_io.checkpoint_save(path.decode(), np.full_like(half, 1e6), 64, 48, 16, 16, 11, frame_fingerprint=hdr.frame_fingerprint ^ 1)
        rc, img, nrays, err = adapter_render(ref, cs, RGBAZ, checker.MODE_FIXED, 11)
        assert rc == 0, err
        assert np.array_equal(img.view(np.uint32), one_shot.view(np.uint32)) and nrays == full_rays
        # ... and so must the checkpoint of the same scene seen from another camera pose
        cs2 = checker.camera_settings(64, 48, (50, 500, 1900), (0, 5, 0), 6.0, (3.6, 2.7), 16)
        rc, img2, _, err = adapter_render(ref, cs2, RGBAZ, checker.MODE_FIXED, 11)
        assert rc == 0, err
        L.adapter_set_checkpoint(None, 0)
        rc, want2, _, err = adapter_render(ref, cs2, RGBAZ, checker.MODE_FIXED, 11)
        assert rc == 0 and np.array_equal(img2.view(np.uint32), want2.view(np.uint32)) and not np.array_equal(img2, one_shot)
    finally:
        L.adapter_set_checkpoint(None, 0)


@pytest.mark.gpu
@needs_adapter
def test_adapter_renders_every_camera_of_the_list():
    """Render takes a camera LIST (integrators.h:14, pathtracer.cpp:210): two cameras of different size and pose in one call, each
    film identical to the one a single-camera call produces (the scene is uploaded and built once)."""
    pos, nrm, uv = scenes.spheres(2, 2, 3, 150.0)
    ref = checker.Ref(pos, nrm, uv)
    cams = [checker.camera_settings(64, 48, (0, 500, 1900), (0, 5, 0), 6.0, (3.6, 2.7), 16),
            checker.camera_settings(40, 56, (300, 400, 1700), (3, -8, 0), 6.0, (2.0, 2.8), 8)]
    singles = []
    for cs in cams:
        rc, img, _, err = adapter_render(ref, cs, RGBAZ, checker.MODE_FIXED, 7)
        assert rc == 0, err
        singles.append(img)
    L = C.CDLL(ADAPTER)
    L.adapter_render_cameras.argtypes = [C.c_void_p, C.POINTER(checker.RefCameraSettings), C.c_int, C.c_int, C.c_uint64, C.POINTER(C.c_void_p),
                                         C.c_char_p, C.c_int]
    arr = (checker.RefCameraSettings * 2)(*cams)
    outs = [np.zeros((cs.width * cs.height, 5), np.float32) for cs in cams]
    ptrs = (C.c_void_p * 2)(*[o.ctypes.data for o in outs])
    err = C.create_string_buffer(512)
    rc = L.adapter_render_cameras(ref.h, arr, 2, checker.MODE_FIXED, 7, ptrs, err, 512)
    assert rc == 0, err.value
    for a, b in zip(outs, singles):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        assert a[:, :3].max() > 0


@pytest.mark.gpu
@needs_adapter
def test_adapter_scene_cache_is_keyed_on_content():
    """ADVICE r1 / VERDICT r1: the adapter used to key its GPU scene on (MeshEngine pointer, face count, texture count). One integrator
    renders, the MeshEngine's vertices are moved IN PLACE (same pointer, same counts), and it renders again: the second film must be the
    moved scene's, i.e. equal to what a fresh engine renders from the moved triangles."""
    from vermilion_b200 import engine
    pos, nrm, uv = scenes.spheres(2, 2, 3, 150.0)
    ref = checker.Ref(pos, nrm, uv)
    cs = checker.camera_settings(64, 48, (0, 500, 1900), (0, 5, 0), 6.0, (3.6, 2.7), 8)
    L = C.CDLL(ADAPTER)
    L.adapter_render_edit_render.argtypes = [C.c_void_p, C.POINTER(checker.RefCameraSettings), C.c_int, C.c_uint64, C.c_float, C.c_void_p, C.c_void_p,
                                             C.c_char_p, C.c_int]
    a, b = np.zeros((64 * 48, 5), np.float32), np.zeros((64 * 48, 5), np.float32)
    err = C.create_string_buffer(512)
    rc = L.adapter_render_edit_render(ref.h, C.byref(cs), checker.MODE_FIXED, 4, 120.0, a.ctypes.data, b.ctypes.data, err, 512)
    assert rc == 0, err.value
    moved = pos.copy()
    moved[:, 1::3] += np.float32(120.0)
    with engine.Engine(device=0) as eng:
        cam = engine.camera_from_settings(64, 48, (0, 500, 1900), (0, 5, 0), 6.0, (3.6, 2.7), 8)
        for p, got in ((pos, a), (moved, b)):
            eng.set_scene(p, nrm, uv)
            eng.set_camera(cam)
            eng.render(mode=engine.MODE_FIXED, seed=4)
            assert np.array_equal(eng.read_image().view(np.uint32), got.view(np.uint32))
    assert not np.array_equal(a, b)


def _adapter_devices(devs, split_samples=False):
    L = C.CDLL(ADAPTER)
    L.adapter_set_devices.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_int]
    L.adapter_set_devices((C.c_int * len(devs))(*devs), len(devs), int(split_samples))


def _gpu_lists():
    import torch
    lists = [[0, 0, 0]]  # one GPU, three members: peer-copy exchange (NCCL refuses duplicate devices)
    if torch.cuda.device_count() >= 2:
        lists.append(list(range(min(torch.cuda.device_count(), 4))))  # NCCL over NVLink
    return lists


@pytest.mark.gpu
@needs_adapter
@pytest.mark.parametrize("mode", [checker.MODE_COMPAT, checker.MODE_FIXED, 2])
def test_adapter_on_several_devices_is_bit_identical_to_one(mode):
    """B200Options::devices: the ONE Render call of the host (renderEngine.cpp:164) spread over several devices by interleaved tiles;
    Camera::mImage is bit-identical to the single-device frame for every estimator (PathTracer compat / fixed, BruteForceTracer)."""
    pos, nrm, uv = scenes.spheres(2, 2, 3, 150.0)
    ref = checker.Ref(pos, nrm, uv)
    cs = checker.camera_settings(200, 136, (0, 500, 1900), (0, 5, 0), 6.0, (3.6, 3.6 * 136 / 200), 16)
    rc, one, rays_one, err = adapter_render(ref, cs, RGBAZ, mode, 11)
    assert rc == 0, err
    try:
        for devs in _gpu_lists():
            _adapter_devices(devs)
            rc, img, rays, err = adapter_render(ref, cs, RGBAZ, mode, 11)
            assert rc == 0, err
            assert np.array_equal(img.view(np.uint32), one.view(np.uint32)), devs
            assert rays == rays_one
            if mode == checker.MODE_FIXED:
                _adapter_devices(devs, split_samples=True)
                rc, img, rays, err = adapter_render(ref, cs, RGBAZ, mode, 11)
                assert rc == 0, err
                assert np.array_equal(img[:, 3:], one[:, 3:]) and np.allclose(img, one, rtol=1e-5, atol=1e-6) and rays == rays_one
    finally:
        _adapter_devices([])

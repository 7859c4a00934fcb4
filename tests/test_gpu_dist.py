"""GPU tests (-m gpu) of the multi-GPU layer behind the C ABI (vmn_dist_*, csrc/dist.cu): one group object, one Render-like call, the
film assembled on the root. On a one-GPU box the group is built over a device list that repeats device 0 — NCCL refuses duplicate
devices, so the group exchanges by peer copies, which runs the same split / pack / gather / unpack code; with two or more GPUs the
same tests run over NCCL (ncclCommInitAll) as well. Reference: the host makes one Render call per frame (renderEngine.cpp:147-166) and
the pixel loop is the parallel axis (pathtracer.cpp:226-227)."""
import numpy as np
import pytest

from vermilion_b200 import engine, scenes

pytestmark = pytest.mark.gpu


def _device_lists():
    import torch
    n = torch.cuda.device_count()
    lists = [[0, 0], [0, 0, 0]]
    if n >= 2:
        lists.append(list(range(min(n, 4))))
    return lists


@pytest.fixture(scope="module")
def scene():
    return scenes.spheres(2, 2, 4, 150.0)


@pytest.fixture(scope="module")
def single(scene):
    pos, nrm, uv = scene
    e = engine.Engine(device=0, max_paths=1 << 20)
    e.set_scene(pos, nrm, uv)
    yield e
    e.close()


def _cam(spp, w=400, h=232):  # 232 is not a multiple of the tile size: the last tile row is partial
    return engine.camera_from_settings(w, h, (0.0, 500.0, 1900.0), (0.0, 0.0, 0.0), 6.0, (3.6, 3.6 * h / w), spp)


@pytest.mark.parametrize("devices", _device_lists(), ids=lambda d: "dev" + "".join(map(str, d)))
@pytest.mark.parametrize("broadcast", [False, True])
def test_group_tiles_reproduce_the_single_gpu_frame_bit_for_bit(scene, single, devices, broadcast):
    pos, nrm, uv = scene
    with engine.DistGroup.local(devices, max_paths=1 << 20) as g:
        assert g.world == len(devices) and g.my_rank == 0
        g.set_scene(pos, nrm, uv, broadcast=broadcast)
        for mode, spp, tile in ((engine.MODE_FIXED, 8, 64), (engine.MODE_COMPAT, 16, 32), (engine.MODE_BRUTEFORCE, 12, 64)):
            cam = _cam(spp)
            g.set_camera(cam)
            single.set_camera(cam)
            g.render(mode=mode, seed=5, split=engine.DIST_SPLIT_TILES, tile=tile)
            single.render(mode=mode, seed=5)
            assert np.array_equal(g.read_image().view(np.uint32), single.read_image().view(np.uint32)), mode
            assert np.array_equal(g.read_image(engine.FILM_SUM), single.read_image(engine.FILM_SUM))
            st = g.stats()
            assert st["world"] == len(devices) and len(st["render_ms"]) == len(devices) and st["rays_traced"] == single.stats()["rays_traced"]


@pytest.mark.parametrize("devices", _device_lists(), ids=lambda d: "dev" + "".join(map(str, d)))
def test_group_sample_slices_sum_to_the_single_gpu_film(scene, single, devices):
    pos, nrm, uv = scene
    with engine.DistGroup.local(devices, max_paths=1 << 20) as g:
        g.set_scene(pos, nrm, uv, broadcast=True)
        cam = _cam(10)  # 10 samples over 3 ranks: slices of 4, 3, 3
        g.set_camera(cam)
        single.set_camera(cam)
        g.render(mode=engine.MODE_FIXED, seed=9, split=engine.DIST_SPLIT_SAMPLES)
        single.render(mode=engine.MODE_FIXED, seed=9)
        a, b = g.read_image(engine.FILM_SUM), single.read_image(engine.FILM_SUM)
        assert np.array_equal(a[:, 3], b[:, 3]) and (a[:, 3] == 10).all()
        assert np.allclose(a, b, rtol=1e-6, atol=1e-6)  # the same samples, added in another order
        assert g.stats()["rays_traced"] == single.stats()["rays_traced"]
        # progressive passes stay local; the pass that asks for it assembles the frame on the root
        g.render(mode=engine.MODE_FIXED, seed=9, split=engine.DIST_SPLIT_SAMPLES, sample_begin=0, sample_end=4)
        g.render(mode=engine.MODE_FIXED, seed=9, split=engine.DIST_SPLIT_SAMPLES, sample_begin=4, sample_end=10, accumulate=True, assemble=True)
        c = g.read_image(engine.FILM_SUM)
        assert np.array_equal(c[:, 3], b[:, 3]) and np.allclose(c, b, rtol=1e-6, atol=1e-6)
        # more ranks than samples: the empty shares contribute zeros
        cam1 = _cam(1)
        g.set_camera(cam1)
        single.set_camera(cam1)
        g.render(mode=engine.MODE_FIXED, seed=9, split=engine.DIST_SPLIT_SAMPLES)
        single.render(mode=engine.MODE_FIXED, seed=9)
        assert np.array_equal(g.read_image(engine.FILM_SUM), single.read_image(engine.FILM_SUM))


def test_group_argument_errors(scene):
    pos, nrm, uv = scene
    with pytest.raises(engine.VmnError):
        engine.DistGroup.local([])
    with engine.DistGroup.local([0, 0], max_paths=1 << 18) as g:
        with pytest.raises(engine.VmnError, match="STATE|camera"):
            g.render()
        g.set_scene(pos, nrm, uv)
        g.set_camera(_cam(8))
        with pytest.raises(engine.VmnError, match="FIXED"):
            g.render(mode=engine.MODE_COMPAT, split=engine.DIST_SPLIT_SAMPLES)
        with pytest.raises(engine.VmnError, match="split"):
            g.render(split=7)
    with engine.DistGroup.local([0]) as g1:  # a world of one needs no NCCL and no exchange
        g1.set_scene(pos, nrm, uv)
        g1.set_camera(_cam(4))
        g1.render(split=engine.DIST_SPLIT_TILES)
        assert (g1.read_image(engine.FILM_SUM)[:, 3] == 4).all()

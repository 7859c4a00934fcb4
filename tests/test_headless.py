"""oracle/_ref/vermilion_headless (vermilion_b200/host/vermilion_headless.cpp): the reference's command line (main.cpp:58-104)
without the Vulkan preview, integrator selectable. `--engine cpu` runs the reference's own PathTracer on an Assimp-loaded OBJ;
`--engine b200` runs the drop-in on the same MeshEngine. The CPU engine seeds from std::random_device (pathtracer.cpp:231), so the
GPU image is judged against the reference's own run-to-run RMSE."""
import json
import os
import subprocess

import numpy as np
import pytest

import checker
from test_film_io import _decode_png
from vermilion_b200 import scenes

HEADLESS = os.path.join(checker.ORACLE_DIR, "_ref", "vermilion_headless")
needs_headless = pytest.mark.skipif(not os.path.exists(HEADLESS), reason="oracle/_ref/vermilion_headless not built")


def run(args, timeout=600):
    r = subprocess.run([HEADLESS] + [str(a) for a in args], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=timeout)
    line = [l for l in r.stdout.splitlines() if l.startswith("{")]
    return r.returncode, (json.loads(line[-1]) if line else None), r.stderr


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"Pf"
        w, h = map(int, f.readline().split())
        assert float(f.readline()) < 0  # little endian
        return np.frombuffer(f.read(), "<f4").reshape(h, w)[::-1]


@needs_headless
def test_cpu_engine_renders_an_obj_like_main_cpp(tmp_path):
    pos, nrm, uv = scenes.cornell()
    obj = scenes.write_obj(str(tmp_path / "cornell.obj"), pos, nrm, uv)
    out = str(tmp_path / "frame")
    rc, info, err = run([obj, "--engine", "cpu", "--res", "48x40", "--spp", 16, "--camera", "0,275,1100,0,0,0", "--film", "6,3.6,3.0",
                         "--out", out, "--quiet"])
    assert rc == 0, err
    assert info["engine"] == "cpu" and info["triangles"] == 30 and (info["width"], info["height"]) == (48, 40)
    img = _decode_png(out + ".png")
    assert img.shape == (40, 48, 4) and (img[:, :, 3] == 255).all()  # alpha = 1 (pathtracer.cpp:321)
    depth = read_pfm(out + "_depth.pfm")  # samples taken: the early-out leaves between sqrt(16)+1 and 16 (pathtracer.cpp:290-311)
    assert depth.shape == (40, 48) and depth.min() >= 5 and depth.max() <= 16
    assert abs(depth.mean() - info["mean_depth_plane"]) < 1e-3
    rc, _, _ = run([str(tmp_path / "missing.obj")])
    assert rc == 3
    rc, _, err = run([])
    assert rc == 1 and "usage" in err


@needs_headless
def test_b200_engine_fails_loudly_without_a_gpu(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    pos, nrm, uv = scenes.cornell()
    obj = scenes.write_obj(str(tmp_path / "cornell.obj"), pos, nrm, uv)
    out = str(tmp_path / "frame")
    rc, info, err = run([obj, "--engine", "b200", "--res", "16x16", "--spp", 16, "--out", out])
    assert rc == 2 and info is None and "no CPU fallback" in err
    assert not os.path.exists(out + ".png")


@pytest.mark.gpu
@needs_headless
@pytest.mark.parametrize("pair", [("cpu", "b200"), ("cpu-bf", "b200-bf")])
def test_b200_engine_matches_the_cpu_engine_within_its_own_noise(tmp_path, pair):
    """Same OBJ, same command line, only --engine differs. The two CPU runs differ by their random_device (PathTracer) or
    time(0)-and-thread-race (BruteForceTracer) seeds; the B200 image must be as close to a CPU run as the CPU runs are to each
    other (x 1.5 + one 8-bit step), and take the same number of samples on average."""
    cpu, gpu = pair
    pos, nrm, uv = scenes.spheres(2, 2, 2, 150.0)  # open to emitter B (meshEngine.cpp:410), 4 x 320 triangles
    obj = scenes.write_obj(str(tmp_path / "spheres.obj"), pos, nrm, uv)
    common = [obj, "--res", "160x120", "--spp", 64, "--camera", "0,500,1900,0,5,0", "--film", "6,3.6,2.7", "--quiet"]
    imgs, infos = {}, {}
    for tag, eng in (("cpu1", cpu), ("cpu2", cpu), ("gpu", gpu)):
        rc, info, err = run(common + ["--engine", eng, "--out", str(tmp_path / tag), "--seed", 5])
        assert rc == 0, err
        imgs[tag] = _decode_png(str(tmp_path / tag) + ".png")[:, :, :3].astype(np.float64) / 255.0
        infos[tag] = info
    rmse = lambda a, b: float(np.sqrt(np.mean((imgs[a] - imgs[b]) ** 2)))
    floor = rmse("cpu1", "cpu2")
    # two BruteForceTracer runs started within the same second share their time(0) seed and can come out almost identical, which
    # would make the floor meaninglessly small: 0.012 is 2.5 x the GPU-vs-CPU RMSE of that integrator at this size (0.0045)
    assert rmse("gpu", "cpu1") <= max(1.5 * floor + 1.0 / 255.0, 0.012), (rmse("gpu", "cpu1"), floor)
    assert abs(imgs["gpu"].mean() - imgs["cpu1"].mean()) <= 0.02
    if cpu == "cpu":  # depth plane = samples taken; BruteForceTracer stores the last hit distance there instead
        assert abs(infos["gpu"]["mean_depth_plane"] - infos["cpu1"]["mean_depth_plane"]) <= 0.05 * infos["cpu1"]["mean_depth_plane"]
    assert infos["gpu"]["rays_traced"] > 160 * 120


@pytest.mark.gpu
@needs_headless
@pytest.mark.parametrize("engine_name", ["b200", "b200-bf"])
def test_headless_on_several_gpus_writes_the_same_png(tmp_path, engine_name):
    """`--gpus N` / `--devices a,b,...`: the one Render call of the command line spread over several devices (B200Options::devices ->
    vmn_dist local group, tiles). The PNG and the depth plane are byte-identical to the single-device run. On a one-GPU box the device list
    repeats device 0 (peer-copy exchange); with two or more GPUs `--gpus 2` runs over NCCL as well."""
    import torch
    pos, nrm, uv = scenes.spheres(2, 2, 2, 150.0)
    obj = scenes.write_obj(str(tmp_path / "spheres.obj"), pos, nrm, uv)
    common = [obj, "--res", "200x136", "--spp", 32, "--camera", "0,500,1900,0,5,0", "--film", "6,3.6,2.448", "--quiet", "--engine", engine_name, "--seed", 5]
    rc, one, err = run(common + ["--out", str(tmp_path / "one")])
    assert rc == 0, err
    variants = [("dup", ["--devices", "0,0,0"])]
    if torch.cuda.device_count() >= 2:
        variants.append(("nccl", ["--gpus", 2]))
    for tag, extra in variants:
        rc, info, err = run(common + extra + ["--out", str(tmp_path / tag)])
        assert rc == 0, err
        assert info["gpus"] == (3 if tag == "dup" else 2) and len(info["render_ms"]) == info["gpus"]
        assert info["rays_traced"] == one["rays_traced"]
        for suffix in (".png", "_depth.pfm"):
            assert open(str(tmp_path / tag) + suffix, "rb").read() == open(str(tmp_path / "one") + suffix, "rb").read(), (tag, suffix)

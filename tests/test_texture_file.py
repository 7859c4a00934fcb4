"""The texture arguments of the reference's command line (main.cpp:71-92) through the reference's OWN MeshEngine::bindTexture
(meshEngine.cpp:74-93). OpenImageIO is absent here; the oracle build's stand-in reads PFM, which is all bindTexture needs
(spec + read_image as float). CPU only: the GPU side of textures is covered by tests/test_gpu_parity.py / test_adapter.py."""
import os

import numpy as np
import pytest

import checker
from common import aimed_rays
from test_headless import HEADLESS, needs_headless, run
from test_film_io import _decode_png
from vermilion_b200 import film_io, scenes

needs_ref = pytest.mark.skipif(not checker.Ref.available("ieee"), reason="oracle/_ref/libvermref.so not built")


def _texture(seed, h, w, scale):
    return np.random.default_rng(seed).random((h, w, 3), dtype=np.float32) * np.float32(scale)


@needs_ref
def test_bind_texture_reads_a_pfm_through_the_reference_loader(tmp_path):
    """A file-bound texture behaves bit for bit like the same texels injected into boundTextures, and like the restatement's."""
    tex = _texture(3, 8, 16, 4.0)
    path = str(tmp_path / "albedo.pfm")
    film_io.write_pfm(path, tex, 16, 8, 3)  # rows top to bottom in, bottom to top in the file
    pos, nrm, uv = scenes.spheres(2, 1, 2, 150.0)  # open to emitter B, so that paths carry light and the texture shows
    from_file, injected, orc = checker.Ref(pos, nrm, uv), checker.Ref(pos, nrm, uv), checker.Oracle(pos, nrm, uv)
    assert from_file.bind_texture_file(path)
    assert not from_file.bind_texture_file(str(tmp_path / "missing.pfm"))  # bindTexture returns false (meshEngine.cpp:78)
    assert not from_file.bind_texture_file(__file__)                       # not an image
    injected.set_texture(tex)
    orc.set_texture(tex)
    box = (np.array([-1900, 10, -1900], np.float32), np.array([1900, 990, 1900], np.float32))
    rays = aimed_rays(pos, 6000, seed=8, spread=0.05, origin_box=box)
    a = from_file.radiance(rays, seed=4, mode=checker.MODE_COMPAT)
    b = injected.radiance(rays, seed=4, mode=checker.MODE_COMPAT)
    c = orc.radiance(rays, seed=4, rng=checker.RNG_MT64, mode=checker.MODE_COMPAT)[0]
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert np.array_equal(a.view(np.uint32), c.view(np.uint32))
    plain = checker.Ref(pos, nrm, uv).radiance(rays, seed=4, mode=checker.MODE_COMPAT)
    assert not np.array_equal(a, plain)  # the texture is on the path (pathtracer.cpp:62-80)


@needs_headless
def test_headless_takes_the_normal_map_and_texture_arguments(tmp_path):
    """vermilion_headless scene normalmap texture --engine cpu-bf: both files are bound with bindTexture (main.cpp:86-92) and
    BruteForceTracer uses them (integrators.cpp:98-106, 141-147), so the frame differs from the untextured one."""
    pos, nrm, uv = scenes.spheres(2, 2, 2, 150.0)
    obj = scenes.write_obj(str(tmp_path / "spheres.obj"), pos, nrm, uv)
    nmap = (_texture(5, 8, 8, 1.0) - np.float32(0.5)) * np.float32(0.6)
    alb = _texture(6, 8, 16, 1.0)
    film_io.write_pfm(str(tmp_path / "n.pfm"), nmap, 8, 8, 3)
    film_io.write_pfm(str(tmp_path / "a.pfm"), alb, 16, 8, 3)
    common = ["--engine", "cpu-bf", "--res", "96x72", "--spp", 16, "--camera", "0,500,1900,0,5,0", "--film", "6,3.6,2.7", "--quiet"]
    rc, info, err = run([obj, str(tmp_path / "n.pfm"), str(tmp_path / "a.pfm")] + common + ["--out", str(tmp_path / "tex")])
    assert rc == 0, err
    rc, info0, err = run([obj] + common + ["--out", str(tmp_path / "plain")])
    assert rc == 0, err
    a = _decode_png(str(tmp_path / "tex.png"))[:, :, :3].astype(np.float64)
    b = _decode_png(str(tmp_path / "plain.png"))[:, :, :3].astype(np.float64)
    assert a.shape == b.shape == (72, 96, 3)
    assert np.sqrt(np.mean((a - b) ** 2)) > 5.0  # far beyond the run-to-run noise of the untextured frame (8-bit units)

"""GPU tests (-m gpu) of the failure paths and of the host-side ordering rules:
  * a tree deeper than the traversal stack is refused at commit (VMN_ERR_STATE), and if one gets through anyway the dropped push is
    REPORTED by the next synchronising call instead of returning a wrong hit silently (VERDICT r1 / ADVICE r1);
  * render_distributed orders the film collective (torch's stream) against an Engine that owns its own non-blocking stream.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

import checker
from vermilion_b200 import distributed as vd
from vermilion_b200 import engine, scenes

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TINY = os.path.join(ROOT, "vermilion_b200", "variants", "libvermilion_b200_tinystack.so")

_CHILD = r"""
import sys, zlib, numpy as np
sys.path[:0] = [%(root)r, %(root)r + "/tests"]
import common
from vermilion_b200 import engine, scenes
pos, nrm, uv = scenes.spheres(2, 1, 5, 150.0)   # 40 960 triangles: seven wide levels
rays = common.aimed_rays(pos, 20000, 3)          # every ray walks down to a leaf
e = engine.Engine(device=0)
try:
    e.set_scene(pos, nrm, uv)
    print("COMMIT_OK depth", e.stats()["wide_depth"], flush=True)
except engine.VmnError as x:
    print("COMMIT_REFUSED", x, flush=True)
    sys.exit(0)
for flags in (0, engine.ISECT_BVH2):
    try:
        h = e.intersect(rays, flags)
        print("INTERSECT_OK", flags, zlib.crc32(h.tobytes()), flush=True)
    except engine.VmnError as x:
        print("INTERSECT_REPORTED", flags, x, flush=True)
"""


def _run_child(extra_env):
    env = dict(os.environ, **extra_env)
    r = subprocess.run([sys.executable, "-c", _CHILD % {"root": ROOT}], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=180)
    return r.stdout


def test_too_deep_tree_is_refused_at_commit_and_overflow_is_reported():
    """The `tinystack` build of the library (3 stack entries, built by __graft_entry__.build()) makes an ordinary scene 'too deep':
    the commit is refused with VMN_ERR_STATE; with the depth check switched off (test hook) every traversal call either REPORTS the
    overflow (VMN_ERR_STATE) or returns exactly what the normal library returns — never a silently wrong hit."""
    assert os.path.exists(TINY), "variants/libvermilion_b200_tinystack.so missing: run __graft_entry__.build()"
    out = _run_child({"VMN_LIB": TINY})
    assert "COMMIT_REFUSED" in out and "STATE" in out and "exceeds the traversal stack" in out, out
    good = _run_child({})
    assert good.count("INTERSECT_OK") == 2, good
    want = {l.split()[1]: l.split()[2] for l in good.splitlines() if l.startswith("INTERSECT_OK")}
    out = _run_child({"VMN_LIB": TINY, "VMN_DEBUG_NO_DEPTH_CHECK": "1"})
    assert "COMMIT_OK" in out, out
    lines = [l for l in out.splitlines() if l.startswith("INTERSECT_")]
    assert len(lines) == 2, out
    for l in lines:
        f = l.split()
        if f[0] == "INTERSECT_OK":
            assert f[2] == want[f[1]], "a traversal with a 3-entry stack returned different hits WITHOUT reporting: " + out
        else:
            assert "traversal stack overflowed" in l, out
    assert any(l.startswith("INTERSECT_REPORTED") for l in lines), out  # seven levels cannot fit three entries


def test_checked_build_runs_every_kernel_without_a_bounds_violation():
    """compute-sanitizer is closed on this GPU pool, so the library carries its own memcheck: the `checked` build (-DVMN_CHECKED=1) tests
    every data-dependent index against its array before the access and raises VMN_DEVERR_BOUNDS -> VMN_ERR_STATE. tools/sanitize_small.py
    drives every kernel of the library on small scenes (and compares the results with the CPU oracle)."""
    lib = os.path.join(ROOT, "vermilion_b200", "variants", "libvermilion_b200_checked.so")
    assert os.path.exists(lib), "variants/libvermilion_b200_checked.so missing: run __graft_entry__.build()"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sanitize_small.py")], env=dict(os.environ, VMN_LIB=lib), stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "all kernels ran" in r.stdout, r.stdout[-3000:]


def test_pipelined_scene_load_equals_the_serial_load(monkeypatch):
    """vmn_scene_set_triangles builds the position-only part of the BVH beside the upload of the normals and uvs, a group broadcasts the
    positions first (r02x/y); VMN_LOAD_OVERLAP=0 is the serial load. Same tree, same film — also when a context is reloaded with a scene of
    another size and back (grow-only staging buffers, the stashed build status)."""
    a = scenes.spheres(2, 1, 4, 150.0)
    b = scenes.spheres(1, 1, 3, 90.0)
    cam = engine.camera_from_settings(320, 200, (0.0, 500.0, 1900.0), (0.0, 0.0, 0.0), 6.0, (3.6, 3.6 * 200 / 320), 8)
    got = {}
    for v in ("1", "0"):
        monkeypatch.setenv("VMN_LOAD_OVERLAP", v)
        with engine.Engine(device=0) as e:
            frames = []
            for scene in (a, b, a):
                e.set_scene(*scene)
                e.set_camera(cam)
                e.render(mode=engine.MODE_FIXED, seed=5)
                st = e.stats()
                frames.append((e.read_image(), st["n_wide_nodes"], st["wide_depth"], st["rays_traced"]))
            assert np.array_equal(frames[0][0], frames[2][0]) and frames[0][1:] == frames[2][1:]
            assert e.stats()["build_ms"] > 0
            got[v] = frames
        with engine.DistGroup.local([0, 0], max_paths=1 << 20) as g:
            g.set_scene(*a, broadcast=True)
            g.set_camera(cam)
            g.render(mode=engine.MODE_FIXED, seed=5, split=engine.DIST_SPLIT_TILES)
            got["g" + v] = g.read_image()
    for fa, fb in zip(got["1"], got["0"]):
        assert np.array_equal(fa[0], fb[0]) and fa[1:] == fb[1:]
    assert np.array_equal(got["g1"], got["g0"]) and np.array_equal(got["g1"], got["1"][0][0])


def test_render_distributed_orders_the_collective_for_a_default_stream_engine():
    import torch
    import torch.distributed as dist
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29541", rank=0, world_size=1, device_id=dev)
    try:
        pos, nrm, uv = scenes.spheres(2, 1, 3, 150.0)
        with engine.Engine(device=0) as e:  # no stream=: the context creates its own non-blocking stream
            e.set_scene(pos, nrm, uv)
            cam = engine.camera_from_settings(640, 360, (0.0, 500.0, 1900.0), (0.0, 0.0, 0.0), 6.0, (3.6, 3.6 * 360 / 640), 16)
            e.set_camera(cam)
            e.render(mode=engine.MODE_FIXED, seed=3)
            want = e.read_image()
            for split in ("sample", "tile"):
                for _ in range(3):  # back-to-back frames: the next render's film memset must not overtake the collective either
                    img = vd.render_distributed(e, dev, engine.MODE_FIXED, 3, split, 0, 1, 16, force_reduce=True)
                    assert np.array_equal(img, want)
            e.render(mode=engine.MODE_BRUTEFORCE, seed=3)
            want_bf = e.read_image()
            for _ in range(3):
                assert np.array_equal(vd.render_bruteforce_distributed(e, dev, 3, 0, 1, force_reduce=True), want_bf)
    finally:
        dist.destroy_process_group()

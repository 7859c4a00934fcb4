"""C1 as BASELINE.json words it — "Cornell box (~30 tris, Assimp-loaded OBJ)": the scene goes through the reference's OWN loading
path, MeshEngine::load -> Assimp::Importer::ReadFile (vendored Assimp 4.1.0, compiled from /root/reference/extern/assimp by
oracle/Makefile) -> processScene -> createBVH (meshEngine.cpp:95-150, 649-756), and what comes out is compared with the CPU
restatement (not gpu) and with the B200 engine behind the Integrator seam (gpu)."""
import numpy as np
import pytest

import checker
from common import aimed_rays
from vermilion_b200 import scenes

needs_ref = pytest.mark.skipif(not checker.Ref.available("ieee"), reason="oracle/_ref/libvermref.so not built")

CORNELL_MATERIALS = [("white", (0.73, 0.73, 0.73), 0, 6), ("red", (0.65, 0.05, 0.05), 6, 8),
                     ("green", (0.12, 0.45, 0.15), 8, 10), ("blocks", (0.73, 0.73, 0.73), 10, 30)]


def cornell_obj(tmp_path):
    pos, nrm, uv = scenes.cornell()
    return scenes.write_obj(str(tmp_path / "cornell.obj"), pos, nrm, uv, CORNELL_MATERIALS), (pos, nrm, uv)


@needs_ref
def test_cornell_obj_survives_the_reference_loader(tmp_path):
    """30 triangles in four `usemtl` groups -> four aiMeshes, walked in createBVH's order: positions and normals come back bit for
    bit (the text is %.9g), v texture coordinates flipped by aiProcess_FlipUVs (meshEngine.cpp:107)."""
    path, (pos, nrm, uv) = cornell_obj(tmp_path)
    ref = checker.Ref.load_file(path)
    assert len(ref.pos) == 30 and ref.n_meshes == 4
    assert np.array_equal(ref.pos, pos.reshape(-1, 9))
    assert np.array_equal(ref.nrm, nrm.reshape(-1, 9))
    flipped = uv.reshape(-1, 3, 2).copy()
    flipped[:, :, 1] = np.float32(1.0) - flipped[:, :, 1]
    assert np.array_equal(ref.uv, flipped.reshape(-1, 6))
    with pytest.raises(RuntimeError):
        checker.Ref.load_file(str(tmp_path / "missing.obj"))  # load() returns false (meshEngine.cpp:132-139)


@needs_ref
def test_loaded_cornell_raycast_and_radiance_match_the_restatement(tmp_path):
    """The reference's MeshEngine holding the Assimp-loaded scene against the oracle port fed with the same triangles: closest-hit
    ids, RayCast records and whole Radiance paths bit-exact."""
    path, _ = cornell_obj(tmp_path)
    ref = checker.Ref.load_file(path)
    orc = checker.Oracle(ref.pos, ref.nrm, ref.uv)
    rays = aimed_rays(ref.pos, 20000, seed=5, spread=0.05)
    a, b = ref.intersect(rays), orc.intersect(rays)
    assert np.array_equal(a["tri"], b["tri"]) and np.array_equal(a["t"].view(np.uint32), b["t"].view(np.uint32))
    ra, rb = ref.raycast(rays), orc.raycast(rays)
    for f in ("P", "dist", "N", "flags", "uv", "Le"):
        assert np.array_equal(ra[f].view(np.uint32), rb[f].view(np.uint32)), f
    pa = ref.radiance(rays[:4000], seed=9, mode=checker.MODE_COMPAT)
    pb = orc.radiance(rays[:4000], seed=9, rng=checker.RNG_MT64, mode=checker.MODE_COMPAT)[0]
    assert np.array_equal(np.asarray(pa).view(np.uint32), np.asarray(pb).view(np.uint32))


@needs_ref
def test_quads_without_normals_are_triangulated_and_smoothed(tmp_path):
    """aiProcess_Triangulate + aiProcess_GenSmoothNormals (meshEngine.cpp:106-109): a cube of six quads, no `vn`, shared corners.
    12 triangles come out, every normal is unit length, and the restatement agrees with the reference on them."""
    p = tmp_path / "cube.obj"
    v = [(x, y, z) for x in (-100.0, 100.0) for y in (200.0, 400.0) for z in (-100.0, 100.0)]
    quads = [(1, 2, 4, 3), (5, 7, 8, 6), (1, 5, 6, 2), (3, 4, 8, 7), (1, 3, 7, 5), (2, 6, 8, 4)]
    p.write_text("".join("v %g %g %g\n" % c for c in v) + "".join("f %d %d %d %d\n" % q for q in quads))
    ref = checker.Ref.load_file(str(p))
    assert len(ref.pos) == 12 and ref.n_meshes == 1
    n = ref.nrm.reshape(-1, 3)
    assert np.allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-5)
    assert not ref.uv.any()  # no texture coordinates: createBVH leaves the uvs alone, the harness walk reads them as 0
    orc = checker.Oracle(ref.pos, ref.nrm, ref.uv)
    rays = aimed_rays(ref.pos, 5000, seed=6, spread=0.1)
    a, b = ref.intersect(rays), orc.intersect(rays)
    assert np.array_equal(a["tri"], b["tri"]) and np.array_equal(a["t"].view(np.uint32), b["t"].view(np.uint32))
    assert (a["tri"] != checker.MISS).mean() > 0.5


@pytest.mark.gpu
@needs_ref
def test_b200_renders_the_assimp_loaded_cornell_through_the_integrator_seam(tmp_path):
    """C1 end to end: OBJ -> the reference's MeshEngine::load -> B200PathTracer::Render called through Integrator* on that
    MeshEngine (the adapter walks sceneMeshes itself) -> the reference Camera's film, against the CPU oracle on the same triangles
    with the shared Philox stream, compat estimator, 16 spp."""
    from test_adapter import ADAPTER, RGBAZ, adapter_render
    import os
    if not os.path.exists(ADAPTER):
        pytest.skip("oracle/_ref/libvermadapter.so not built")
    path, _ = cornell_obj(tmp_path)
    ref = checker.Ref.load_file(path)
    cs = checker.camera_settings(96, 96, (0, 275, 1100), (0, 0, 0), 6.0, (3.6, 3.6), 16)
    rc, img, nrays, err = adapter_render(ref, cs, RGBAZ, checker.MODE_COMPAT, 21)
    assert rc == 0, err
    oimg, onrays = checker.Oracle(ref.pos, ref.nrm, ref.uv).render(ref.camera_fields(cs), checker.MODE_COMPAT, checker.RNG_PHILOX, 21)
    assert np.array_equal(img[:, 3:], oimg[:, 3:])  # alpha and the samples-taken plane
    assert (img[:, :3].view(np.uint32) == oimg[:, :3].view(np.uint32)).all(1).mean() > 0.97
    assert abs(nrays - onrays) <= 0.002 * onrays


@needs_ref
@pytest.mark.parametrize("with_nrm,with_uv", [(False, False), (False, True), (True, False)])
def test_obj_writer_face_forms_load(tmp_path, with_nrm, with_uv):
    """The four OBJ corner forms (v, v/vt, v//vn, v/vt/vn): positions always survive bit for bit and in order; missing normals are
    generated by aiProcess_GenSmoothNormals (unit length, smoothed across corners at the same position), missing uvs stay zero."""
    pos, nrm, uv = scenes.spheres(1, 1, 1, 150.0)  # one 80-triangle icosphere
    path = scenes.write_obj(str(tmp_path / "s.obj"), pos, nrm if with_nrm else None, uv if with_uv else None)
    ref = checker.Ref.load_file(path)
    assert np.array_equal(ref.pos, pos.reshape(-1, 9))
    n = ref.nrm.reshape(-1, 3, 3)
    assert np.allclose(np.linalg.norm(n, axis=2), 1.0, atol=1e-5)
    if with_nrm:
        assert np.array_equal(ref.nrm, nrm.reshape(-1, 9))
    else:
        # corners at the same position are smoothed together even though the OBJ shares no indices (the step welds by position):
        # on an icosphere that is close to the radial direction the generator itself uses
        assert ((n * nrm.reshape(-1, 3, 3)).sum(2) > 0.98).all()
    if with_uv:
        assert np.array_equal(ref.uv.reshape(-1, 3, 2)[:, :, 0], uv.reshape(-1, 3, 2)[:, :, 0])
    else:
        assert not ref.uv.any()

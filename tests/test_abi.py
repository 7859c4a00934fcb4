"""CPU tests: the C-ABI library loads without a GPU, exports exactly what include/vermilion_b200.h declares, struct layouts of the
ctypes mirror match the header, and the product path fails loudly (no CPU fallback) when no CUDA device is present."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from vermilion_b200 import engine

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "vermilion_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vmn_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = engine.load_library()
    decl = declared_functions()
    assert len(decl) >= 20
    for name in decl:
        assert hasattr(lib, name), f"{name} declared in include/vermilion_b200.h but not exported"
    assert sorted(engine.EXPORTS) == decl
    assert lib.vmn_abi_version() == 2


def test_no_other_symbols_leak():
    out = subprocess.run(["nm", "-D", "--defined-only", engine.LIB_PATH], stdout=subprocess.PIPE, text=True).stdout
    names = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert names == set(declared_functions()), names ^ set(declared_functions())


def test_struct_layouts_match_header(tmp_path):
    """Compile a C probe against the header and compare sizeof/offsetof with the ctypes and numpy mirrors."""
    probe = tmp_path / "probe.c"
    probe.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "vermilion_b200.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                     "sizeof(vmn_config),sizeof(vmn_ray),sizeof(vmn_hit),sizeof(vmn_raycast_rec),sizeof(vmn_sphere),sizeof(vmn_camera),"
                     "sizeof(vmn_render_params),sizeof(vmn_stats),offsetof(vmn_render_params,tile_size),offsetof(vmn_stats,kernel_launches));return 0;}\n")
    exe = tmp_path / "probe"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(probe), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], stdout=subprocess.PIPE, text=True, check=True).stdout.split()]
    want = [C.sizeof(engine.Config), engine.RAY.itemsize, engine.HIT.itemsize, engine.RAYCAST.itemsize, engine.SPHERE.itemsize,
            C.sizeof(engine.Camera), C.sizeof(engine.RenderParams), C.sizeof(engine.Stats), engine.RenderParams.tile_size.offset,
            engine.Stats.kernel_launches.offset]
    assert got == want


def test_reference_spheres_table():
    s = engine.reference_spheres()  # pure host code: meshEngine.cpp:377-500
    assert len(s) == 8 and s["radius"][0] == np.float32(3.5) and s["radius"][1] == 250
    assert (s["radius"][2:] == np.float32(5e7)).all()
    assert np.allclose(s["emission"][0], (0, 7.5, 15)) and np.allclose(s["emission"][1], 15.2)
    assert list(s["flags"]) == [1, 0, 0, 0, 1, 1, 1, 0]


def test_fails_loudly_without_cuda_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(engine.VmnError, match="no CUDA device|CUDA"):
        engine.Engine(device=0)


def test_static_scratch_entry_point_rejects_a_null_context_without_touching_the_gpu():
    """vmn_intersect_static (no reference counterpart; include/vermilion_b200.h) checks its context before anything else."""
    lib = engine.load_library()
    assert lib.vmn_intersect_static(None, None, 0, None, 0) == 1  # VMN_ERR_INVALID
    src = open(HEADER).read()
    assert "#define VMN_STATIC_MAX_RAYS 65536u" in src and engine.STATIC_MAX_RAYS == 65536

import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _alias_golden_generator():
    """tests/golden/make_golden.py's scene()/rays_for() under the module name tests_golden_scenes"""
    import importlib.util
    spec = importlib.util.spec_from_file_location("tests_golden_scenes", os.path.join(ROOT, "tests", "golden", "make_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["tests_golden_scenes"] = mod
    spec.loader.exec_module(mod)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Native libraries are built in-tree by __graft_entry__.build(); build them if a fresh checkout lacks them."""
    need = [os.path.join(ROOT, "vermilion_b200", "libvermilion_b200.so"), os.path.join(ROOT, "vermilion_b200", "host", "libvermilion_host.so"),
            os.path.join(ROOT, "oracle", "liboracle.so")]
    if not all(os.path.exists(p) for p in need):
        import __graft_entry__
        __graft_entry__.build()
    _alias_golden_generator()

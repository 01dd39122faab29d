import importlib.util
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

PKG_DIR = os.path.join(ROOT, "ray-tracing-in-one-weekend_b200")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_pkg():
    """Import the hyphen-named package directory as module `rt_b200`."""
    if "rt_b200" in sys.modules:
        return sys.modules["rt_b200"]
    spec = importlib.util.spec_from_file_location(
        "rt_b200", os.path.join(PKG_DIR, "__init__.py"), submodule_search_locations=[PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["rt_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def rt():
    return load_pkg()


@pytest.fixture(scope="session")
def orc():
    import oracle_lib
    oracle_lib.lib()
    return oracle_lib


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


@pytest.fixture(scope="session")
def cover_rows():
    return golden("cover_small.npz")["spheres"]


@pytest.fixture(scope="session")
def three_rows():
    return golden("three_small.npz")["spheres"]


def cam_for(rt_mod, scene: str, W: int, H: int):
    import oracle_lib
    p = oracle_lib.COVER_CAM if scene == "cover" else oracle_lib.THREE_CAM
    return rt_mod.Camera(lookfrom=p["lookfrom"], lookat=p["lookat"], vup=p["vup"], vfov=p["vfov"],
                         aspect_ratio=W / H, aperture=p["aperture"], focus_dist=p["focus"])

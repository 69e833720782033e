import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _ensure_built():
    import __graft_entry__ as g
    need = [g.LIB, os.path.join(ROOT, "oracle", "librt_oracle.so"), os.path.join(ROOT, "tests", "hostsim", "libhostsim.so")]
    if not all(os.path.exists(p) for p in need):
        g.build()


@pytest.fixture(scope="session")
def rt():
    _ensure_built()
    import rtpkg
    return rtpkg.load()


@pytest.fixture(scope="session")
def lib(rt):
    return rt.load_library()


@pytest.fixture(scope="session")
def orc():
    _ensure_built()
    import oracle_py
    return oracle_py


@pytest.fixture(scope="session")
def oracle_kind(orc):
    """The strongest oracle available: the reference's own compiled sources, else the port."""
    return "reference" if orc.available("reference") else "port"


@pytest.fixture(scope="session")
def sim(rt):
    import hostsim_py
    return hostsim_py


@pytest.fixture(scope="session")
def gpu(lib):
    from my_raytracer_b200 import capi
    st = lib.rt_init_device(0)
    if st != 0:
        pytest.fail("gpu test without a usable CUDA device: " + lib.rt_last_error().decode())
    return capi


GOLDEN = os.path.join(ROOT, "tests", "golden")

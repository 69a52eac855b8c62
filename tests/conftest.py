"""pytest configuration: the `gpu` marker, native builds, and scene / oracle fixtures shared by the suite."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def _build_once():
    """libpsgpu.so, the C oracle and the host-emulation helper are built on demand (prebuilt files travel to the GPU box)."""
    need = [os.path.join(ROOT, "parsip_b200", "libpsgpu.so"), os.path.join(ROOT, "oracle", "libps_oracle.so")]
    if not all(os.path.exists(p) for p in need):
        sys.path.insert(0, ROOT)
        import __graft_entry__ as g
        g.build()
    emu = os.path.join(ROOT, "tests", "hostemu", "libhostemu.so")
    if not os.path.exists(emu):
        subprocess.check_call(["make", "-f", os.path.join(ROOT, "tests", "hostemu", "Makefile")])
    fac = os.path.join(ROOT, "tests", "facade", "libfacade_probe.so")
    if not os.path.exists(fac):
        subprocess.check_call(["make", "-f", os.path.join(ROOT, "tests", "facade", "Makefile")])


_build_once()


def has_gpu() -> bool:
    import parsip_b200 as ps
    try:
        return ps.load_library().psgpu_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # a -m gpu run on a box without a device must fail loudly, not skip: there is no CPU fallback to hide behind
    pass


@pytest.fixture(scope="session")
def ps():
    import parsip_b200
    return parsip_b200


@pytest.fixture(scope="session")
def ctx(ps):
    c = ps.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="session")
def rng():
    return np.random.RandomState(1234)

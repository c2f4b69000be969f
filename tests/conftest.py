import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module("real-time-gpu-ray-tracer_b200")


@pytest.fixture(scope="session")
def oc():
    import oracle_ctypes
    return oracle_ctypes


@pytest.fixture(scope="session")
def port(oc):
    return oc.Oracle("port")


@pytest.fixture(scope="session")
def ref(oc):
    if not os.path.exists(oc.REF_PATH):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return oc.Oracle("ref")


def parity_scenes(S):
    """Small versions of the BASELINE configs used by both the CPU and the GPU parity tests."""
    a = S.c1a(frame=0, width=160, height=90, spp=4, depth=4)
    b = S.c1a(frame=37, width=96, height=54, spp=4, depth=6)
    b.focus_disk_radius = 0.5
    b.name = "C1a-frame37-defocus"
    return [a, b, S.c1b(width=128, height=72, spp=4, depth=4),
            S.c2(width=160, height=90, spp=4, depth=10, nx=100, nz=50),
            S.c5(n_instances=200, n_meshes=4, width=96, height=54, spp=1, depth=4, subdiv=1),
            S.c3(os.path.join(ROOT, "tests", "golden", "particle_frame0.vtk"), frame=12, width=160, height=90, spp=1, depth=10)]


N_PARITY = 6

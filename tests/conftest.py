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


def edge_scenes(S):
    """More ground for pinning the oracle port to the reference (tests/golden/edge*.npz): every primitive type with both
    materials in one scene, a one-instance mesh under a rotated / non-uniformly scaled transform (what the path kernel's SINGLE
    instantiation runs on), the config-4 generator, an odd window with defocus late in the demo animation, and the quirks Q7
    (sample count floored to a square), Q6 (depth 1) and Q14 (singular instance transform) run through the reference itself."""
    a = S.c1_mixed()                      # its glass (an extension the reference lacks) becomes metal: all types x both reference materials
    from importlib import import_module
    abi = import_module(S.__name__.rsplit(".", 1)[0] + ".abi")
    for arr in (a.spheres, a.parallelograms, a.triangles):
        glass = arr["material_type"] == abi.DIELECTRIC
        arr["material_index"][glass] = arr["material_index"][glass] % len(a.metals)
        arr["material_type"][glass] = abi.METAL
    a.dielectrics = a.dielectrics[:0]
    a.name = "C1-mixed-rough-metal"
    b = S.c2(width=96, height=54, spp=4, depth=8, nx=40, nz=20)
    b.xforms[0]["shift"] = (0.5, -0.25, 1.0); b.xforms[0]["rotate_deg"] = (20.0, 35.0, -10.0); b.xforms[0]["scale"] = (1.5, 0.75, 1.2)
    b.name = "C2-transformed-instance"
    c = S.c4(width=128, height=72, spp=1, depth=10, nx=80, nz=16)
    d = S.c1a(frame=200, width=75, height=41, spp=9, depth=5)
    d.focus_disk_radius = 0.25
    d.name = "C1a-frame200-odd-window-defocus"
    e = S.c1a(frame=3, width=48, height=27, spp=5, depth=1)
    e.xforms[1]["scale"] = 0.0
    e.name = "C1a-quirks-Q6-Q7-Q14"
    return [a, b, c, d, e]


N_EDGE = 5

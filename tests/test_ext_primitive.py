"""SURVEY 8f row 4 — "add a geometry type" as code, not prose: the reference's procedure is ten steps in ten files
(files/AddGeometryType.txt:1-35); here a type is one row of the primitive table + one PrimOps specialisation
(csrc/rt_prims.h).  The proof is a FOURTH type, a disc, that exists only in tests/ext/rt_prim_disc.h and in test variants of
the libraries (librtb200_disc.so, libemul_disc.so); the oracle port has its own, independently written disc
(oracle_port.cpp ORACLE_EXT_DISC).  No reference counterpart: the comparison pins the plumbing (build, wide BVH, traversal,
shading, instancing, materials all reach the new type through the table), not reference arithmetic."""
import os

import numpy as np
import pytest

import emul_ctypes as ec

SEED = 0x5EED
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DISC_LIB = os.path.join(ROOT, "real-time-gpu-ray-tracer_b200", "librtb200_disc.so")
disc_dt = np.dtype([("c", np.float32, 3), ("r", np.float32), ("n", np.float32, 3), ("d", np.float32),
                    ("material_type", np.uint32), ("material_index", np.uint32), ("_p", np.uint32, 2)])
DISC = 3


def disc_scene(pkg, width=160, height=90, spp=4, depth=6):
    """C1b's ground and a third of its spheres, plus 40 discs of all three materials: single-disc instances (some rotated and
    scaled) and one instance of 16 discs (a multi-primitive BLAS of the new type)."""
    S, abi = pkg.scenes, pkg.abi
    rs = np.random.RandomState(7)
    sc = S.c1b(width=width, height=height, spp=spp, depth=depth)
    sc.name = "C1b+discs"
    sc.center, sc.target, sc.fov = (0.0, 3.0, 8.0), (0.0, 0.8, 0.0), 75.0
    keep = 33
    sc.spheres = sc.spheres[:keep].copy()
    sc.dielectrics = np.zeros(1, abi.dielectric_dt)
    sc.dielectrics[0] = ((1.0, 1.0, 1.0), 1.5)
    n = 40
    d = np.zeros(n, disc_dt)
    for k in range(n):
        nv = rs.normal(size=3)
        nv[1] = abs(nv[1]) + 0.3
        nv = (nv / np.linalg.norm(nv)).astype(np.float32)
        nv = nv / np.float32(np.sqrt(np.float32(nv[0] * nv[0] + nv[1] * nv[1]) + np.float32(nv[2] * nv[2])))     # unit in float32
        c = np.array([rs.uniform(-6, 6), rs.uniform(0.4, 2.5), rs.uniform(-4, 4)], np.float32)
        dd = np.float32(nv[0] * c[0])
        dd = np.float32(dd + np.float32(nv[1] * c[1]))
        dd = np.float32(dd + np.float32(nv[2] * c[2]))
        mt = [abi.ROUGH, abi.METAL, abi.DIELECTRIC][k % 3]
        mi = 0 if mt == abi.DIELECTRIC else int(rs.randint(0, len(sc.roughs) if mt == abi.ROUGH else len(sc.metals)))
        d[k] = (c, rs.uniform(0.3, 0.9), nv, dd, mt, mi, (0, 0))
    sc.ext_records = d
    kinds = [(abi.PARALLELOGRAM, 0)] + [(abi.SPHERE, i) for i in range(keep)] + [(DISC, i) for i in range(24)]
    ins = S.single_instances(kinds)
    group = np.zeros(1, abi.instance_desc_dt)
    group[0]["type"], group[0]["first_prim"], group[0]["prim_count"] = DISC, 24, 16
    sc.instances = np.concatenate([ins, group])
    sc.xforms = S.identity_xforms(len(sc.instances))
    for k in range(1 + keep, 1 + keep + 24, 3):          # every third single disc: rotated, scaled, shifted
        sc.xforms[k]["rotate_deg"] = tuple(rs.uniform(0, 360, 3))
        sc.xforms[k]["scale"] = tuple(rs.uniform(0.6, 1.6, 3))
        sc.xforms[k]["shift"] = tuple(rs.uniform(-1, 1, 3))
    sc.xforms[-1]["shift"] = (0.0, 0.5, -1.0)
    sc.xforms[-1]["rotate_deg"] = (0.0, 30.0, 0.0)
    return sc


def against_port(pkg, oc, sc, got):
    h = oc.Oracle("port_disc").scene(sc, 21)
    img, rad, rays = h.render(SEED)
    t, inst, prim = h.primary(SEED)
    same = (got["inst"] == inst) & (got["prim"] == prim)
    assert ((prim >> 28) == DISC).sum() > 200, "the discs are not visible"
    assert same.mean() >= 0.9999
    hit = same & (inst >= 0)
    assert (np.abs(got["t"][hit] - t[hit]) / np.abs(t[hit])).max() <= 1e-5
    assert float(np.sqrt(np.mean((got["radiance"] - rad) ** 2))) <= 1e-3
    print(f"{sc.name}: ids {same.mean():.6f}, bit-exact radiance {np.array_equal(got['radiance'], rad)}, rays {got['rays']} / {rays}")


def test_twin_with_a_fourth_type_matches_the_port(pkg, oc):
    """`not gpu`: the product's cores compiled by g++ with the disc in the table, against the port's own disc."""
    sc = disc_scene(pkg)
    e = ec.Emul(sc, variant="disc")
    r = e.render(SEED)
    assert r["overflow"] == 0
    against_port(pkg, oc, sc, r)
    # the standard twin does not know the type: the scene is refused, not mis-rendered
    with pytest.raises(AssertionError):
        ec.Emul(sc)


@pytest.mark.gpu
def test_gpu_with_a_fourth_type(pkg, oc):
    """The CUDA path of the disc variant (every kernel reaches the type through the table: bounds, build, collapse, path
    kernel, wave and one-thread-per-pixel comparison kernels) against the port's disc and, bit for bit, against the twin."""
    sc = disc_scene(pkg)
    twin = ec.Emul(sc, variant="disc").render(SEED)
    for flags in (0, pkg.abi.FLAG_WAVEFRONT, pkg.abi.FLAG_MEGAKERNEL, pkg.abi.FLAG_SAH_BUILD):
        r = pkg.Renderer(device=0, flags=flags | pkg.abi.FLAG_RECORD_PRIMARY_AOV, lib_path=DISC_LIB)
        try:
            assert r.lib.rt_primitive_type_count() == 4
            r.commitScene(sc)
            out = r.renderFrame(0, SEED)
            H, W = sc.height, sc.width
            got = dict(rays=out.rays, radiance=r.readFramebuffer(pkg.abi.BUF_RADIANCE).reshape(H, W, 3).copy(),
                       rgba8=r.readFramebuffer(pkg.abi.BUF_RGBA8).reshape(H, W, 4).copy(),
                       t=r.readFramebuffer(pkg.abi.BUF_PRIMARY_T).reshape(H, W).copy(),
                       inst=r.readFramebuffer(pkg.abi.BUF_PRIMARY_INSTANCE).reshape(H, W).copy(),
                       prim=r.readFramebuffer(pkg.abi.BUF_PRIMARY_PRIM).reshape(H, W).copy())
        finally:
            r.cleanup()
        assert out.stack_overflow == 0
        against_port(pkg, oc, sc, got)
        assert got["rays"] == twin["rays"] and np.array_equal(got["radiance"], twin["radiance"]) and np.array_equal(got["rgba8"], twin["rgba8"])
    # the product build has three types: the commit is refused with RT_ERR_ARGUMENT
    r = pkg.Renderer(device=0)
    try:
        assert r.lib.rt_primitive_type_count() == 3
        with pytest.raises(pkg.RtError) as e:
            r.commitExtensionPrimitives(sc.ext_records)
        assert e.value.code == -2
    finally:
        r.cleanup()

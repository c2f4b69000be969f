#!/usr/bin/env python
"""Generates tests/golden/*.npz by running THE REFERENCE ITSELF (its own sources under /root/reference compiled as
C++ into oracle/_ref/libref_oracle.so by oracle/Makefile) on the small parity scenes.  Run in the build container
(where /root/reference exists); the GPU box only reads the committed .npz files.

    python tests/golden/make_golden.py

Each file holds, for one scene of tests/conftest.py::parity_scenes (scene*.npz) or ::edge_scenes (edge*.npz): the RGBA8 frame written by the reference's own
`render` kernel body (src/Global/Kernel.cu:105-147) through the surf2Dwrite shim, the float radiance and ray count
of the rayColor loop, the primary-hit AOVs (t, instance, primitive) recorded by the id-recording copy of
TLAS::hit/BLAS::hit, the derived camera, the instance matrices/boxes, and a few closest-hit probes.
xorwow.npz holds curand_uniform streams for a few (seed, subsequence) pairs.
"""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import oracle_ctypes as oc  # noqa: E402
from conftest import edge_scenes, parity_scenes  # noqa: E402

SEED = 0x5EED
BUILD_SEED = 20261104


def probe_rays(sc, n=512, seed=7):
    rs = np.random.RandomState(seed)
    o = np.array(sc.center, np.float32) + rs.uniform(-1, 1, (n, 3)).astype(np.float32)
    tgt = np.array(sc.target, np.float32) + rs.uniform(-6, 6, (n, 3)).astype(np.float32)
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([o, d.astype(np.float32)], axis=1).astype(np.float32)


def main():
    pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
    ref = oc.Oracle("ref")
    sets = [(f"scene{k}", sc) for k, sc in enumerate(parity_scenes(pkg.scenes))] + [(f"edge{k}", sc) for k, sc in enumerate(edge_scenes(pkg.scenes))]
    for stem, sc in sets:
        h = ref.scene(sc, BUILD_SEED)
        rgba_kernel = h.render_kernel(SEED)
        rgba, rad, rays = h.render(SEED)
        assert np.array_equal(rgba, rgba_kernel), "ref driver's rayColor loop differs from the reference render kernel"
        t, inst, prim = h.primary(SEED)
        pr = probe_rays(sc)
        pt, pi, pp, pn = h.trace(pr)
        np.savez_compressed(os.path.join(HERE, f"{stem}.npz"), name=sc.name, seed=SEED, rgba8=rgba, radiance=rad, rays=np.uint64(rays),
                            t=t, inst=inst, prim=prim, camera=h.camera(), instances=h.instances(),
                            probe_rays=pr, probe_t=pt, probe_inst=pi, probe_prim=pp, probe_normal=pn)
        print(f"{stem}.npz  {sc.name}: rays={rays} hit-pixels={(inst >= 0).mean():.3f}")
        h.close()
    pairs = [(0, 0), (0x5EED, 0), (0x5EED ^ 1234, 1234), (0x5EED ^ 57599, 57599), (2 ** 40 + 17, 2 ** 20 + 3), (12345678901234567, 2073599)]
    np.savez_compressed(os.path.join(HERE, "xorwow.npz"), pairs=np.array(pairs, np.uint64),
                        uniforms=np.stack([ref.uniforms(s, q, 64) for s, q in pairs]))
    print("xorwow.npz")


if __name__ == "__main__":
    main()

"""Pixel indices beyond 2^21 (frames of more than 2 M pixels take the hex-digit jump tables for the top index bits): a
32 x 140000 frame at 1 spp, rendered on the GPU, against the CPU oracle on scan-lines whose pixel indices exceed 2^21."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import oracle_ctypes as oc  # noqa: E402

pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
SEED = 0x5EED
sc = pkg.scenes.c1a(frame=0, width=32, height=140000, spp=1, depth=4)          # the centre rows (the scene) have indices just above 2^21
lines = [65535, 65536, 69900, 69990, 70000, 70040]    # pixel index = 32 * y + x: 2^21 starts at y = 65536
h = oc.Oracle("port").scene(sc, 7)
want = {y: h.render(SEED, y, y + 1, rgba8=False)[1][y] for y in lines}
h.close()
if "--oracle-only" in sys.argv:
    print("oracle ok", {y: float(want[y].sum()) for y in lines})
    sys.exit(0)
r = pkg.Renderer(device=0)
r.commitScene(sc)
r.renderFrame(0, SEED)
rad = r.readFramebuffer(pkg.abi.BUF_RADIANCE).reshape(sc.height, sc.width, 3)
r.cleanup()
for y in lines:
    assert np.array_equal(rad[y], want[y]), f"line {y} differs"
print("high pixel indices ok")

import sys, importlib, json, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import oracle_ctypes as oc
pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
# small first: compare image against ours
sc = pkg.scenes.c2(width=320, height=180, spp=4, depth=10, nx=100, nz=50)
ms, host, img = oc.ref_cuda_bench(sc, 1, 2)
r = pkg.Renderer(0); r.commitScene(sc); r.renderFrame(0, 0x5EED)
ours = r.readFramebuffer(pkg.abi.BUF_RGBA8).reshape(180, 320, 4); r.cleanup()
d = np.abs(ours[..., :3].astype(int) - img[..., :3].astype(int))
print("small: ref ms", ms, host, "mean |diff| (different RNG seeds, 4 spp):", d.mean(), "mean ours/ref", ours[..., :3].mean(), img[..., :3].mean())
sc = pkg.scenes.c2(width=1920, height=1080, spp=int(sys.argv[1]) if len(sys.argv) > 1 else 64, depth=10)
ms, host, img = oc.ref_cuda_bench(sc, 1, 2)
print("C2 full: ref frame ms", ms, host)

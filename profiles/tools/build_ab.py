"""A/B of the collapse kernel's grid size (RTB200_COLLAPSE_BLOCKS_PER_SM is read once per process: run once per value):
BLAS build of config 2 (1 M triangles) and of a 500-particle config-3 batch, best of 5."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
S = pkg.scenes
r = pkg.Renderer(device=0)
sc = S.c2(width=64, height=36, spp=1)
r.commitGeometryData(sc.spheres, sc.parallelograms, sc.triangles)
r.commitMaterialData(sc.roughs, sc.metals)
r.configureInstances(sc.instances, sc.xforms)
t = [r.buildAccelerationStructure().blas_ms for _ in range(6)][1:]
print("bps", os.environ.get("RTB200_COLLAPSE_BLOCKS_PER_SM", "default"), "c2 blas ms best %.3f median %.3f" % (min(t), sorted(t)[len(t) // 2]))
r.cleanup()
sc = S.c5(width=64, height=36, spp=1)
r = pkg.Renderer(device=0)
r.commitGeometryData(sc.spheres, sc.parallelograms, sc.triangles)
r.commitMaterialData(sc.roughs, sc.metals)
r.configureInstances(sc.instances, sc.xforms)
t = [r.buildAccelerationStructure() for _ in range(6)][1:]
print("   c5 blas ms best %.3f tlas best %.3f" % (min(x.blas_ms for x in t), min(x.tlas_ms for x in t)))
r.cleanup()

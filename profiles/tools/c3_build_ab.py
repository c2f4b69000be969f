"""Config 3's rebuild (500 particle BLASes, 629 k triangles, one batched LBVH) under the collapse variants: the environment
variables are read once per process, so run once per setting.  Prints lbvh_rebuild_ms / tlas_ms / render_ms (bench.c3_rebuild)."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
d = bench.c3_rebuild(pkg, 0)
tag = " ".join("%s=%s" % (k[7:], v) for k, v in sorted(os.environ.items()) if k.startswith("RTB200_COLLAPSE") or k == "RTB200_LEAF_MAX") or "default"
print("%-48s rebuild %.3f ms  tlas %.3f  render %.3f  launches %d" % (tag, d["lbvh_rebuild_ms"], d["tlas_ms"], d["render_ms_1080p_1spp"], d["build_kernel_launches"]))

"""Critical path of the headline frame ("first commit the evidence").

A pixel is sequential work (its XORWOW stream is consumed in order: sample after sample, bounce after bounce), so a
synchronous frame cannot end before its longest pixel has traced all of its rays one after the other.

  part A  (CPU, `--histogram`): rays per pixel of the C2 frame (1920x1080, 64 spp, depth 10) on every 8th scan-line, with
          the CPU twin (tests/emul: the product's own per-pixel function compiled for the host, bit-identical paths) ->
          histogram, maximum, and the same for the tiny probe frame of part B.  Run in the build container.
  part B  (GPU, `--latency`): the product renders the SAME scene at 8x4 pixels (one warp on an otherwise idle GPU) at
          64 spp / depth 10 and at 1 spp; kernel time / rays of the longest pixel = latency of one ray when nothing hides
          it.  Uses only the product (no twin, no oracle) and part A's JSON for the per-pixel ray counts of the probe.

Output: profiles/r02_tail_evidence.json (part A) and gpurun_out/r02_tail_latency.json (part B).
"""
import argparse
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
SEED = 0x5EED
PROBE = (8, 4)


def histogram(out_path, stride):
    import emul_ctypes as ec
    pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
    sc = pkg.scenes.c2(spp=64)
    e = ec.Emul(sc)
    rows = list(range(stride // 2, sc.height, stride))
    per = np.concatenate([e.pixel_rays(SEED, y, y + 1) for y in rows], axis=0)
    spp = 64
    edges = [spp * k for k in range(1, 12)]                       # 64 = every sample missed, 640 = every path ran to depth 10
    hist, _ = np.histogram(per, bins=edges)
    sc_p = pkg.scenes.c2(spp=64, width=PROBE[0], height=PROBE[1])
    probe = ec.Emul(sc_p).pixel_rays(SEED, 0, PROBE[1])
    sc_p1 = pkg.scenes.c2(spp=1, width=PROBE[0], height=PROBE[1])
    probe1 = ec.Emul(sc_p1).pixel_rays(SEED, 0, PROBE[1])
    rec = {
        "what": "rays per pixel, C2 (1 000 000 triangles, 1920x1080, 64 spp, depth 10, seed 0x5EED), CPU twin, every %d-th scan-line" % stride,
        "rows_sampled": len(rows), "pixels_sampled": int(per.size),
        "rays_per_pixel": {"min": int(per.min()), "mean": float(per.mean()), "median": float(np.median(per)),
                           "p90": float(np.percentile(per, 90)), "p99": float(np.percentile(per, 99)), "max": int(per.max())},
        "histogram": {"bin_edges_rays": edges, "pixels": [int(v) for v in hist]},
        "frame_rays_estimate": float(per.mean()) * sc.width * sc.height,
        "probe_8x4_64spp": {"rays_per_pixel": probe.astype(int).ravel().tolist(), "max": int(probe.max()), "total": int(probe.sum())},
        "probe_8x4_1spp": {"rays_per_pixel": probe1.astype(int).ravel().tolist(), "max": int(probe1.max()), "total": int(probe1.sum())},
    }
    with open(out_path, "w") as f:
        json.dump(rec, f, indent=1)
    print(json.dumps(rec["rays_per_pixel"]), rec["probe_8x4_64spp"]["max"], rec["probe_8x4_1spp"]["max"])


def latency(hist_path, out_path):
    pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
    rec = json.load(open(hist_path))
    res = {}
    for tag, spp in (("probe_8x4_64spp", 64), ("probe_8x4_1spp", 1)):
        sc = pkg.scenes.c2(spp=spp, width=PROBE[0], height=PROBE[1])
        r = pkg.Renderer(device=0)
        try:
            r.commitScene(sc)
            ms = []
            for k in range(12):
                out = r.renderFrame(k, SEED)
                ms.append(float(out.trace_ms))                          # the path kernel alone (CUDA events)
            rays = int(out.rays)
        finally:
            r.cleanup()
        best = min(ms[2:])
        longest = rec[tag]["max"]
        assert rays == rec[tag]["total"], (rays, rec[tag]["total"])      # the same paths as the twin counted
        res[tag] = {"kernel_ms_best": best, "kernel_ms_median": float(np.median(ms[2:])), "rays": rays, "longest_pixel_rays": longest,
                    "us_per_ray_on_the_longest_pixel": 1e3 * best / longest}
    lat = res["probe_8x4_64spp"]["us_per_ray_on_the_longest_pixel"]
    mx = rec["rays_per_pixel"]["max"]
    res["critical_path"] = {"longest_pixel_rays_c2": mx, "idle_us_per_ray": lat, "ms": 1e-3 * lat * mx,
                            "what": "lower bound of ONE synchronous launch of the frame or of any shard that holds the longest pixel: its rays, one after the other, at the idle-GPU latency"}
    with open(out_path, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--histogram", action="store_true")
    ap.add_argument("--latency", action="store_true")
    ap.add_argument("--stride", type=int, default=8)
    a = ap.parse_args()
    hp = os.path.join(ROOT, "profiles", "r02_tail_evidence.json")
    if a.histogram:
        histogram(hp, a.stride)
    if a.latency:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        latency(hp, os.path.join(ROOT, "gpurun_out", "r02_tail_latency.json"))

#!/usr/bin/env python
"""Trace-driven model of the path kernel's warp votes (no GPU needed).

The CPU twin (tests/emul) walks every ray of a band of the config-2 frame exactly as persist8_kernel's SINGLE instantiation does and
writes, per pixel, the sequence of steps its lane takes: NODE (one wide-node test), LEAF (one primitive test), SHADE (ends the ray,
starts the next; takes the next pixel when the pixel is done).  emul_simt_run then replays the kernel's vote (kernels_persist.cu:
node while >= node_min lanes stand on a node, else shade once >= shade_min lanes wait, else leaf, else node, else shade) over warps of
32 lanes that share one pixel queue in tile order, and counts the steps of every phase and the lanes that took part.

1. Validation: lanes per phase against the ncu capture of the same kernel (profiles/r02s1_persist8_c2_hotlines.txt), and the measured
   threshold sweep (profiles/r02s_sweep.txt, two-level kernel) against the model's cost of the same thresholds.
2. Calibration: warp instructions per step of each phase = the capture's instruction share of the phase x its 113.0 G warp
   instructions / the model's step count.
3. Forecast of step-scheduling variants, at the issue utilisation the kernel has today (0.66).

Usage: python profiles/tools/simt_model.py [--bands 9] [--rows 8] [--warps-per-band 24] > profiles/r02_simt_model.txt
"""
import argparse
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import emul_ctypes as ec  # noqa: E402

SEED = 0x5EED
# the capture (profiles/r02s1_persist8_c2_raw_extract.json, _hotlines.txt classified by source line)
WARP_INST_PER_FRAME = 113.02e9
RAYS_PER_FRAME = 298382765
SHARE = {"node": 0.507, "leaf": 0.177 + 0.015, "shade": 0.148 + 0.014, "iteration": 0.064 + 0.064 + 0.008}   # advance + vote + rest
MEASURED_LANES = {"node": 20.3, "leaf": 10.9, "shade": 11.2}
MEASURED_MRAYS = 2123.0
KEYS = ("node_steps", "leaf_steps", "shade_steps", "node_lanes", "leaf_lanes", "shade_lanes", "leaf_rounds", "leaf_lane_rounds", "iterations", "tokens")


def add(a, b):
    return {k: a.get(k, 0) + b[k] for k in KEYS}


def run(bands, warps, node_min, shade_min, variant):
    tot = {}
    for e in bands:
        tot = add(tot, e.simt_run(warps, node_min, shade_min, variant))
    return tot


def cost(r, c, leaf_overhead=1.0):
    return c["node"] * r["node_steps"] + c["leaf"] * leaf_overhead * r["leaf_rounds"] + c["shade"] * r["shade_steps"] + c["iteration"] * r["iterations"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bands", type=int, default=9)
    ap.add_argument("--rows", type=int, default=8)
    ap.add_argument("--warps-per-band", type=int, default=24, help="1920 x 8 pixels over 24 warps = 20 pixels per lane (the frame: 18)")
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    pkg = importlib.import_module("real-time-gpu-ray-tracer_b200")
    sc = pkg.scenes.c2()
    H = sc.height
    y0s = [int((k + 0.5) * H / args.bands) // 4 * 4 for k in range(args.bands)]
    out = {"bands": [(y, y + args.rows) for y in y0s], "warps_per_band": args.warps_per_band}
    t0 = time.time()

    def make(two_level):
        hs = []
        for y in y0s:
            e = ec.Emul(sc)
            e.simt_tokens(SEED, y, y + args.rows, two_level=two_level)
            hs.append(e)
        return hs

    single = make(False)
    base = run(single, args.warps_per_band, 18, 10, 0)
    rays = base["shade_lanes"]
    print(f"config 2, 1920x1080, 64 spp, depth 10; {args.bands} bands of {args.rows} rows at y = {y0s}; {rays} rays, {base['tokens']} steps; "
          f"{time.time() - t0:.0f} s of CPU")
    print(f"per ray: {base['node_lanes'] / rays:.2f} node steps, {base['leaf_lanes'] / rays:.2f} primitive tests (frame, measured on the GPU: 13.92 / 4.31)")
    print("\n1. lanes per step, model vs ncu (SINGLE instantiation, thresholds 18 / 10)")
    lanes = {"node": base["node_lanes"] / base["node_steps"], "leaf": base["leaf_lanes"] / base["leaf_steps"], "shade": base["shade_lanes"] / base["shade_steps"]}
    for k in ("node", "leaf", "shade"):
        print(f"   {k:6s} model {lanes[k]:5.1f}   ncu {MEASURED_LANES[k]:5.1f}")
    out["lanes_model"], out["lanes_ncu"] = lanes, MEASURED_LANES

    # 2. calibration: warp instructions per step
    scale = RAYS_PER_FRAME / rays
    steps_frame = {"node": base["node_steps"] * scale, "leaf": base["leaf_rounds"] * scale, "shade": base["shade_steps"] * scale, "iteration": base["iterations"] * scale}
    c = {k: SHARE[k] * WARP_INST_PER_FRAME / steps_frame[k] for k in SHARE}
    print("\n2. warp instructions per step (capture's instruction share x 113.0 G / model's steps per frame)")
    for k in c:
        print(f"   {k:10s} {c[k]:7.1f}   ({steps_frame[k] / 1e6:8.1f} M steps per frame, {100 * SHARE[k]:.1f} % of the instructions)")
    out["warp_inst_per_step"] = c
    base_cost = cost(base, c)
    weighted = (c["node"] * base["node_lanes"] + c["leaf"] * base["leaf_lanes"] + c["shade"] * base["shade_lanes"] + c["iteration"] * 32 * base["iterations"]) / base_cost
    print(f"   lanes per warp instruction implied by the model: {weighted:.1f} (ncu: 16.15; the steps themselves diverge inside, which the model does not see)")

    # 3. validation on the measured threshold sweep (two-level kernel: a TLAS-root node step and an instance entry in front of every ray)
    print("\n3. threshold sweep of the two-level kernel: measured Mrays/s (profiles/r02s_sweep.txt) vs the model's cost, both relative to 20 / 8")
    two = make(True)
    ref_m = None
    sweep = [((20, 8), 1902.6), ((16, 8), 1871.9), ((18, 8), 1903.2), ((22, 8), 1877.0), ((24, 8), 1811.7),
             ((20, 4), 1716.9), ((20, 6), 1838.9), ((20, 10), 1922.8), ((20, 12), 1891.4), ((20, 16), 1827.2), ((18, 10), 1928.0), ((22, 6), 1826.6)]
    ref_cost = cost(run(two, args.warps_per_band, 20, 8, 0), c)
    rows = []
    for (nm, sm), mr in sweep:
        r = run(two, args.warps_per_band, nm, sm, 0)
        rel_model = ref_cost / cost(r, c)
        rows.append({"node_min": nm, "shade_min": sm, "measured_rel": mr / 1902.6, "model_rel": rel_model})
        print(f"   node_min {nm:2d} shade_min {sm:2d}   measured {mr / 1902.6:6.3f}   model {rel_model:6.3f}")
    out["sweep"] = rows
    # SINGLE against two-level at the thresholds in use
    two_base = cost(run(two, args.warps_per_band, 18, 10, 0), c)
    print(f"   two-level -> SINGLE at 18 / 10: model x{two_base / base_cost:.3f}, measured x{2123.0 / 1931.0:.3f}")
    out["single_vs_two_level"] = {"model": two_base / base_cost, "measured": 2123.0 / 1931.0}
    del two

    # 4. forecasts
    print("\n4. forecasts for the SINGLE kernel (relative issue cost -> Mrays/s at today's issue utilisation; today = 2123)")
    fc = []

    def line(name, r, leaf_overhead=1.0, note=""):
        rel = base_cost / cost(r, c, leaf_overhead)
        nl = r["node_lanes"] / r["node_steps"]
        ll = r["leaf_lane_rounds"] / max(1, r["leaf_rounds"])
        sl = r["shade_lanes"] / r["shade_steps"]
        fc.append({"variant": name, "relative_speed": rel, "mrays": MEASURED_MRAYS * rel, "node_lanes": nl, "leaf_lanes_per_round": ll, "shade_lanes": sl})
        print(f"   {name:64s} x{rel:5.3f} -> {MEASURED_MRAYS * rel:5.0f}   lanes N {nl:4.1f} L {ll:4.1f} S {sl:4.1f}  {note}")

    line("today (18 / 10, one primitive per lane and LEAF step)", base)
    r2 = run(single, args.warps_per_band, 18, 10, 2)
    line("all pending primitives of a lane in one LEAF step (RT_LEAF_LOOP)", r2, note="measured on the GPU: -15 % (with 52-88 B of spills)")
    r1 = run(single, args.warps_per_band, 18, 10, 1)
    for ov in (1.0, 1.25, 1.5):
        line(f"pending primitives of the WARP handed out to its 32 lanes, leaf round x{ov:.2f}", r1, ov)
    for nm, sm in ((16, 10), (18, 8), (18, 12), (20, 10), (22, 12), (24, 16)):
        line(f"thresholds {nm} / {sm}", run(single, args.warps_per_band, nm, sm, 0))
    for nm, sm in ((20, 12), (22, 12), (24, 14)):
        line(f"warp-shared leaf step (x1.25) with thresholds {nm} / {sm}", run(single, args.warps_per_band, nm, sm, 1), 1.25)
    out["forecasts"] = fc
    ideal = (c["node"] * base["node_lanes"] + c["leaf"] * base["leaf_lanes"] + c["shade"] * base["shade_lanes"]) / 32.0
    print(f"   all 32 lanes busy in every step, no vote overhead (bound of any scheduling of this instruction mix): x{base_cost / ideal:5.3f} -> {MEASURED_MRAYS * base_cost / ideal:5.0f}")
    out["all_lanes_bound"] = base_cost / ideal
    if args.json:
        json.dump(out, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()

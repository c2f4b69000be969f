"""`not gpu`: the JSON lines bench.py printed on the B200 (committed under profiles/) carry every key the bench contract
names, with consistent values; and bench.py's argument surface is the one the driver uses."""
import glob
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE = ["metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data",
        "config", "e2e", "gpu_launches", "clocks"]


def lines(pattern):
    out = []
    for p in sorted(glob.glob(os.path.join(ROOT, "profiles", pattern))):
        text = open(p).read().strip()
        try:
            d = json.loads(text)                           # pretty-printed copy of the line
        except json.JSONDecodeError:
            d = json.loads(text.splitlines()[-1])          # the line as printed
        out.append((os.path.basename(p), d))
    return out


@pytest.mark.parametrize("name,d", lines("r01[i-n]_bench_*.json"))
def test_committed_lines_follow_the_contract(name, d):
    for k in BASE:
        assert k in d, f"{name}: missing {k}"
    assert d["unit"] == "Mrays/s" and d["higher_is_better"] is True and d["data"] == "synthetic" and d["dtype"] == "f32"
    assert d["vs_baseline"] is None                      # BASELINE.md holds no published number for this metric
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["steps"] >= 1 and d["warmup"] >= 1
    assert "workload" in d["config"] and "model" not in d["config"]
    e = d["e2e"]
    assert e["value"] > 0 and e["unit"] == "Mrays/s" and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert d["gpu_launches"] > 0
    c = d["clocks"]
    assert c["sm_mhz"] > 0.9 * c["sm_max_mhz"]
    assert not set(c["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    if "roofline" in d:
        r = d["roofline"]
        assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
        assert "persist8_kernel" in r["kernel"] or "wave" in r["kernel"] or "render_pixels" in r["kernel"]
    if "cpu_baseline" in d:
        b = d["cpu_baseline"]
        assert b["kind"] in ("reference", "port") and b["cores"] >= 1 and b["value"] > 0 and b["sample"]


def test_headline_line_has_roofline_and_cpu_baseline():
    d = dict(lines("r01l_bench_n1.json"))["r01l_bench_n1.json"]
    assert d["n_gpus"] == 1 and "configs[1]" in d["config"]["workload"]
    assert d["roofline"]["traffic"] and d["roofline"]["l2"]["peak"] > d["roofline"]["peak"]
    assert d["cpu_baseline"]["kind"] == "reference" and d["reference_cuda"]["value"] > 0
    assert d["value"] / d["reference_cuda"]["value"] > 2.0          # north_star: at least 2x the reference's own CUDA build
    assert d["c3_particle_rebuild"]["lbvh_rebuild_ms"] <= 1.0       # north_star: LBVH rebuild of a particle frame within 1 ms


def test_bench_argument_surface():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True, check=True).stdout
    for flag in ("--gpus", "--steps", "--warmup", "--impl", "--config"):
        assert flag in out

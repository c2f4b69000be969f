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


@pytest.mark.parametrize("name,d", lines("r02[a-z]_bench_*.json"))
def test_round2_lines_follow_the_contract(name, d):
    """Round 2: the roofline names what binds (L2 when the measured DRAM traffic / the working set say so), the frame loop
    (frames in flight) is described beside the throughput, `parallelism` sits outside `config`."""
    if d.get("impl") == "reference":
        # the reference arm: the reference's own C++ on the host cores, same metric / config, no GPU work
        assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
        assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and "configs[1]" in d["config"]["workload"]
        return
    for k in BASE:
        assert k in d, f"{name}: missing {k}"
    assert d["unit"] == "Mrays/s" and d["higher_is_better"] is True and d["data"] == "synthetic" and d["dtype"] == "f32"
    assert d["vs_baseline"] is None and d["value"] > 0 and d["ms_per_step"] > 0
    assert "workload" in d["config"] and "model" not in d["config"]
    if name != "r02b_bench_c3_n1.json":                       # printed before the animation line moved the key as well
        assert "parallelism" not in d["config"] and "parallelism" in d
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert d["gpu_launches"] > 0
    c = d["clocks"]
    assert c["sm_mhz"] > 0.9 * c["sm_max_mhz"] and not set(c["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    if "frame_loop" in d:
        f = d["frame_loop"]
        assert 1 <= f["frames_per_window"] <= 8
        if f["frames_per_window"] > 1:
            assert f["max_frame_latency_ms"] >= f["synchronous_frame_ms"] * 0.5
    if "roofline" in d:
        r = d["roofline"]
        assert r["bound"] in ("l2", "hbm") and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
        assert abs(r["achieved"] - r["algorithmic_bytes_per_launch"] / (r["kernel_ms"] * 1e-3) / 1e9) < 1e-6 * r["achieved"]
        if r["bound"] == "l2":
            assert r["working_set_bytes"] < 126e6 or (r["traffic"] is not None and r["traffic"] < 0.1 * r["algorithmic_bytes_per_launch"])
            assert r["hbm"]["peak"] > 5000 and r["peak"] > r["hbm"]["peak"]
    if "reference_cuda" in d and "value" in d["reference_cuda"]:
        rc = d["reference_cuda"]
        assert rc["spp"] == d["config"]["spp"] and rc["builds"] >= 1 and rc["min"] <= rc["median"] <= rc["max"]
        assert d["vs_reference_cuda"] > 2.0                    # north_star: at least 2x the reference's own CUDA build


def test_round2_scaling_record():
    """The frames-in-flight loop on 1/N shards of config 2 (bench.py --emulate-shard N, one GPU) and, where measured, on N GPUs
    with the frame ring: the per-GPU rate stays above 85 % of the N = 1 rate (north_star)."""
    got = dict(lines("r02b_bench_*.json"))
    one = got["r02b_bench_n1_short.json"]["ms_per_step"]
    for n in (2, 4, 8):
        shard = got[f"r02b_bench_shard{n}.json"]
        assert one / (n * shard["ms_per_step"]) >= 0.85, f"1/{n} shard"
        sync = shard["frame_loop"]["synchronous_frame_ms"]
        assert shard["ms_per_step"] < sync                      # the window beats one frame at a time
    real = dict(lines("r02d_bench_n*.json"))
    if "r02d_bench_n1.json" in real:
        base = real["r02d_bench_n1.json"]["value"]
        for name, d in real.items():
            n = d["n_gpus"]
            if n > 1:
                assert "frame ring" in d["parallelism"]
                assert d["value"] / base / n >= 0.85, f"{name}: {d['value'] / base / n:.3f} of the single-GPU rate per GPU"
                assert abs(d["e2e"]["value"] / d["value"] - 1.0) < 0.05


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

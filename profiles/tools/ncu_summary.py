#!/usr/bin/env python
"""Summarise one kernel of an ncu report (--set full) as JSON: duration, DRAM/L2 traffic, issue/ALU utilisation,
lane efficiency, stall mix.  Usage: ncu_summary.py report.ncu-rep kernel-regex out.json [traffic_latest.json config [spp]]

With the optional arguments the DRAM bytes per launch are also recorded in traffic_latest.json under `config` ("c2", "c4",
...) together with the hash of the kernel's sources (tools/kernel_hash.py); bench.py reads that into roofline.traffic and
drops it when the sources no longer hash the same.  `spp`: the capture ran at that many samples per pixel (bench.py
scales the bytes to the workload's sample count)."""
import csv, io, json, subprocess, sys

rep, pat, out = sys.argv[1], sys.argv[2], sys.argv[3]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "-k", "regex:" + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units, vals = rows[0], rows[1], rows[2]
m = {h: (v, u) for h, u, v in zip(hdr, units, vals)}


def num(k, scale=None):
    v, u = m[k]
    x = float(v.replace(",", ""))
    if scale:
        x *= scale.get(u, 1.0)
    return x


BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
TIME = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
stalls = {k.split("issue_stalled_")[1].split("_per_issue")[0]: float(v[0]) for k, v in m.items()
          if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio")}
tot = sum(stalls.values()) or 1.0
s = {
    "kernel": m["Kernel Name"][0], "grid": m["Grid Size"][0], "block": m["Block Size"][0],
    "duration_ms": num("gpu__time_duration.sum", TIME),
    "dram_bytes_read": num("dram__bytes_read.sum", BYTES), "dram_bytes_write": num("dram__bytes_write.sum", BYTES),
    "l2_throughput_pct": num("lts__throughput.avg.pct_of_peak_sustained_elapsed"),
    "l2_hit_rate_pct": num("lts__t_sector_hit_rate.pct"), "l1_hit_rate_pct": num("l1tex__t_sector_hit_rate.pct"),
    "issue_active_pct": num("sm__issue_active.avg.pct_of_peak_sustained_elapsed"),
    "alu_pipe_pct": num("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    "warps_active_pct": num("sm__warps_active.avg.pct_of_peak_sustained_active"),
    "threads_per_warp_instruction": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "registers_per_thread": m.get("launch__registers_per_thread", ("?", ""))[0],
    "stall_share": {k: round(v / tot, 4) for k, v in sorted(stalls.items(), key=lambda kv: -kv[1]) if v / tot >= 0.005},
}
s["dram_bytes_per_launch"] = s["dram_bytes_read"] + s["dram_bytes_write"]
json.dump(s, open(out, "w"), indent=1)
if len(sys.argv) > 5:
    import os
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from kernel_hash import kernel_source_hash
    try:
        allrec = json.load(open(sys.argv[4]))
        if "dram_bytes_per_launch" in allrec:
            allrec = {}
    except Exception:
        allrec = {}
    rec = {"dram_bytes_per_launch": s["dram_bytes_per_launch"], "kernel": s["kernel"], "from": out, "kernel_source_sha256": kernel_source_hash()}
    if len(sys.argv) > 6:
        rec["scale_to_spp"] = int(sys.argv[6])
    allrec[sys.argv[5]] = rec
    json.dump(allrec, open(sys.argv[4], "w"), indent=1)
print(json.dumps(s, indent=1))

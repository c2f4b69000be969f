#!/usr/bin/env python
"""Registers / spills / shared memory of every kernel in librtb200.so, from the `-Xptxas -v` logs the csrc Makefile keeps beside the
objects (*.o.ptxas.log).  No GPU needed.  Usage: resource_usage.py > profiles/<tag>_resource_usage.txt"""
import glob, os, re, subprocess

CSRC = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "real-time-gpu-ray-tracer_b200", "csrc")
rows = []
for log in sorted(glob.glob(os.path.join(CSRC, "*.o.ptxas.log"))):
    name = None
    spill = (0, 0, 0)
    for line in open(log, errors="replace"):
        m = re.search(r"Compiling entry function '(\S+)' for 'sm_100a'", line)
        if m:
            name = m.group(1)
            continue
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
        if m:
            spill = tuple(int(x) for x in m.groups())
            continue
        m = re.search(r"Used (\d+) registers(?:, used (\d+) barriers)?(?:, (\d+) bytes cumulative stack size)?(?:, (\d+) bytes smem)?", line)
        if m and name:
            dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
            dem = re.sub(r"\(anonymous namespace\)::", "", dem)
            dem = re.sub(r"\(.*", "", dem).replace("void ", "")
            rows.append((os.path.basename(log).replace(".o.ptxas.log", ".cu"), dem, int(m.group(1)), spill[0], spill[1], spill[2], int(m.group(4) or 0)))
            name = None
print(f"{'file':22s} {'kernel':52s} {'regs':>4s} {'stack':>5s} {'spill st':>8s} {'spill ld':>8s} {'smem B':>7s}")
for r in rows:
    print(f"{r[0]:22s} {r[1]:52s} {r[2]:4d} {r[3]:5d} {r[4]:8d} {r[5]:8d} {r[6]:7d}")

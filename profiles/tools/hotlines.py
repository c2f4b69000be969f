#!/usr/bin/env python
"""Summarise an ncu report's source page by CUDA source line: share of stall samples, of executed warp instructions,
average active threads, top stall reason.  Usage: hotlines.py report.ncu-rep [kernel-regex] [top-N]"""
import csv, io, subprocess, sys

rep = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 else None
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"]
if pat:
    cmd += ["-k", "regex:" + pat]
txt = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, cur_file, lines, seen = None, None, [], set()
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        if (cur_file, r[1]) in seen:
            break           # first captured launch only
        seen.add((cur_file, r[1]))
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < 10:
        continue
    if r[0] != "":          # a CUDA source line carrying the aggregate of its SASS
        d = dict(zip(hdr[2:], r[2:]))
        d["line"], d["src"], d["file"] = r[0], r[1], cur_file
        lines.append(d)

def num(d, k):
    try:
        return float(d.get(k, "0").replace(",", "") or 0)
    except ValueError:
        return 0.0
tot_s = sum(num(d, "# Samples") for d in lines) or 1
tot_i = sum(num(d, "Instructions Executed") for d in lines) or 1
tot_t = sum(num(d, "Thread Instructions Executed") for d in lines)
stalls = [k for k in hdr if k.startswith("stall_") and "Not Issued" not in k]
print(f"samples={tot_s:.0f} warp-inst={tot_i:.0f} thread-inst={tot_t:.0f} avg-threads={tot_t / tot_i:.2f}")
agg = {}
for d in lines:
    for k in stalls:
        agg[k] = agg.get(k, 0) + num(d, k)
print("stall mix: " + "  ".join(f"{k[6:]}={100 * v / tot_s:.1f}%" for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
lines.sort(key=lambda d: -num(d, "# Samples"))
for d in lines[:top]:
    s = max(stalls, key=lambda k: num(d, k))
    print(f"{100 * num(d, '# Samples') / tot_s:5.1f}%smp {100 * num(d, 'Instructions Executed') / tot_i:5.1f}%inst thr={num(d, 'Thread Instructions Executed') / max(1.0, num(d, 'Instructions Executed')):4.1f} "
          f"{s[6:]:>10s} {d['file']}:{d['line']:>4s} {d['src'].strip()[:110]}")

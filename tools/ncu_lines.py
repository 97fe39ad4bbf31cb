#!/usr/bin/env python
"""Per-source-line hot spots of one kernel from an .ncu-rep (needs -lineinfo builds).
usage: ncu_lines.py report.ncu-rep [top N] [kernel-name regex] [lo-hi line ranges to sum ...]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
kern = ["-k", "regex:" + sys.argv[3]] if len(sys.argv) > 3 else []
out = subprocess.run(["ncu", "-i", rep, *kern, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
data = []
fname = ""
hdr = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if len(r) > 8 and r[0] == "Line No":
        hdr = r
        ii = hdr.index("Instructions Executed")
        ss = hdr.index("# Samples")
        continue
    if hdr and len(r) > 8 and r[0].isdigit():
        try:
            data.append((int(r[ii]), int(r[ss]), fname, int(r[0]), r[1].strip()[:100]))
        except ValueError:
            pass
tot = sum(d[0] for d in data) or 1
tots = sum(d[1] for d in data) or 1
print(f"total warp instructions {tot}, samples {tots}")
print(" inst%  smpl%  file:line  source")
for d in sorted(data, reverse=True)[:top]:
    print(f"{d[0] / tot * 100:5.1f}% {d[1] / tots * 100:5.1f}%  {d[2]}:{d[3]}  {d[4]}")
for rng in sys.argv[4:]:
    name, lo, hi = rng.split(":")[0], *map(int, rng.split(":")[1].split("-"))
    ins = sum(d[0] for d in data if lo <= d[3] <= hi and d[2].startswith(name))
    smp = sum(d[1] for d in data if lo <= d[3] <= hi and d[2].startswith(name))
    print(f"{rng}: {ins / tot * 100:5.1f}% instr {smp / tots * 100:5.1f}% samples")

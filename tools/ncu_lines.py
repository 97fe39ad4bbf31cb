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
        stall_cols = [(k, h[6:]) for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        continue
    if hdr and len(r) > 8 and r[0].isdigit():
        try:
            st = sorted(((int(r[k] or 0), n) for k, n in stall_cols), reverse=True)[:3]
            data.append((int(r[ii]), int(r[ss]), fname, int(r[0]), r[1].strip()[:70],
                         " ".join(f"{n}:{v}" for v, n in st if v)))
        except ValueError:
            pass
tot = sum(d[0] for d in data) or 1
tots = sum(d[1] for d in data) or 1
print(f"total warp instructions {tot}, samples {tots}")
print(" inst%  smpl%  file:line  source")
by_samples = "--by-samples" in sys.argv
if by_samples:
    sys.argv.remove("--by-samples")
for d in sorted(data, key=lambda d: d[1] if by_samples else d[0], reverse=True)[:top]:
    print(f"{d[0] / tot * 100:5.1f}% {d[1] / tots * 100:5.1f}%  {d[2]}:{d[3]}  {d[4]}   [{d[5]}]")
for rng in sys.argv[4:]:
    name, lo, hi = rng.split(":")[0], *map(int, rng.split(":")[1].split("-"))
    ins = sum(d[0] for d in data if lo <= d[3] <= hi and d[2].startswith(name))
    smp = sum(d[1] for d in data if lo <= d[3] <= hi and d[2].startswith(name))
    print(f"{rng}: {ins / tot * 100:5.1f}% instr {smp / tots * 100:5.1f}% samples")

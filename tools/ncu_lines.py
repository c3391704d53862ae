#!/usr/bin/env python
"""ncu_lines.py <report.ncu-rep> [top] -- stall samples and executed instructions per CUDA source line (needs -lineinfo)."""
import csv
import subprocess
import sys

def num(x):
    try:
        return int(x.replace(",", ""))
    except ValueError:
        return 0


rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows, fname, hdr, out = list(csv.reader(raw.splitlines())), None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0].isdigit():
        d = dict(zip(hdr[4:], r[4:]))
        out.append((fname, int(r[0]), r[1].strip(), num(d["# Samples"]), num(d["Instructions Executed"]), d))
ts, ti = sum(o[3] for o in out), sum(o[4] for o in out)
print(f"total samples {ts}, warp instructions {ti}")
keys = [k for k in hdr[4:] if k.startswith("stall_") and "Not Issued" not in k]
for f, ln, src, s, i, d in sorted(out, key=lambda o: -o[3])[:top]:
    st = sorted(((num(d[k]), k[6:]) for k in keys), reverse=True)[:2]
    print(f"{f[:14]:14s} {ln:4d} {100 * s / ts:5.1f}% smp {100 * i / ti:5.1f}% ins  {st[0][1]}:{st[0][0]} {st[1][1]}:{st[1][0]}  | {src[:90]}")

#!/usr/bin/env python
"""Per-source-line view of an .ncu-rep captured with --import-source on: samples and executed
warp-instructions per CUDA source line, hottest first.  usage: ncu_lines.py rep [top]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur_file, hdr, rows = None, None, []
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] in ("File Path", "File Name"):
        cur_file, hdr = r[1].split("/")[-1], None
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr and r[0].isdigit() and len(r) >= len(hdr) - 2 and r[2] == "-":
        d = {}
        for k, v in zip(hdr, r):
            d.setdefault(k, v)
        try:
            rows.append((int(d.get("# Samples", "0") or 0), int(d.get("Instructions Executed", "0") or 0),
                         cur_file, int(r[0]), r[1].strip()[:90]))
        except ValueError:
            pass
ts, ti = sum(x[0] for x in rows) or 1, sum(x[1] for x in rows) or 1
print("total samples %d, warp-instructions %d" % (ts, ti))
for s, i, f, ln, src in sorted(rows, reverse=True)[:top]:
    print("%5.1f%% smp %5.1f%% inst  %s:%d  %s" % (100.0 * s / ts, 100.0 * i / ti, f, ln, src))

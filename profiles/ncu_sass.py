#!/usr/bin/env python
"""Per-SASS-instruction view of an .ncu-rep captured with --import-source on: samples, executed
warp-instructions and the stall reasons sampled at each instruction of a window around a pattern.
usage: ncu_sass.py rep kernel-substring sass-pattern [before] [after]   (run here, no GPU)"""
import collections
import csv
import io
import subprocess
import sys

rep, ksub, pat = sys.argv[1], sys.argv[2], sys.argv[3]
before = int(sys.argv[4]) if len(sys.argv) > 4 else 80
after = int(sys.argv[5]) if len(sys.argv) > 5 else 5
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
hdr, fn, sel = None, None, []
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "Kernel Name":
        fn = r[1]
    elif r[0] == "Address":
        hdr = r
    elif hdr and ksub in (fn or "") and r[0].startswith("0x"):
        sel.append(r)
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
i_s, i_i = hdr.index("# Samples"), hdr.index("Instructions Executed")
total = sum(int(r[i_s] or 0) for r in sel)
hits = [k for k, r in enumerate(sel) if pat in r[1]]
k0 = hits[-1] if hits else 0
tot, smp, ins = collections.Counter(), 0, 0
for r in sel[max(0, k0 - before):k0 + after]:
    st = {hdr[i][6:]: int(r[i]) for i in stall if r[i] not in ("", "0", "-")}
    for a, b in st.items():
        tot[a] += b
    smp += int(r[i_s] or 0)
    ins += int(r[i_i] or 0)
    print(r[1][:60].ljust(60), r[i_s].rjust(4), r[i_i].rjust(8), st)
print("window: %d samples of %d in the kernel (%.1f %%), %d warp-instructions; stalls: %s"
      % (smp, total, 100.0 * smp / max(total, 1), ins, dict(tot.most_common())))

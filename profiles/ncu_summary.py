#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key raw metrics, stall mix, hottest SASS.
usage: python profiles/ncu_summary.py gpurun_out/prof.ncu-rep [kernel-index]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_dynamic",
        "launch__shared_mem_per_block_static", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__cycles_elapsed.avg",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "lts__t_sectors_op_write.sum",
        "lts__t_sectors_op_read.sum", "sm__inst_executed_pipe_fp64.sum",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"]
for w in want:
    if w in hdr:
        i = hdr.index(w)
        print("%-64s %s %s" % (w, rows[1][i], [r[i] for r in rows[2:]]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
kern, cur = [], None
for r in csv.reader(io.StringIO(src)):
    if r and r[0] == "Kernel Name":
        cur = {"hdr": None, "rows": []}
        kern.append(cur)
        continue
    if cur is None:
        continue
    if cur["hdr"] is None:
        cur["hdr"] = r
        continue
    cur["rows"].append(r)
k = kern[which]
h, R = k["hdr"], k["rows"]
si, ie = h.index("# Samples"), h.index("Instructions Executed")
stalls = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
tot = {h[i]: 0 for i in stalls}
ts = 0
for r in R:
    ts += int(r[si])
    for i in stalls:
        tot[h[i]] += int(r[i])
print("\nSASS instructions: %d   warp-instructions executed: %d   samples: %d" %
      (len(R), sum(int(r[ie]) for r in R), ts))
for n, v in sorted(tot.items(), key=lambda x: -x[1])[:9]:
    print("  %-24s %6d  %5.1f%%" % (n, v, 100.0 * v / max(ts, 1)))
print("\ninstructions / samples per 100-instruction block of SASS:")
for a in range(0, len(R), 100):
    seg = R[a:a + 100]
    print("  %5d  inst %9d  samples %6d   %s" % (a, sum(int(r[ie]) for r in seg),
                                                  sum(int(r[si]) for r in seg), seg[0][1].strip()[:48]))
print("\nhottest SASS by samples:")
for i in sorted(sorted(range(len(R)), key=lambda i: -int(R[i][si]))[:36]):
    r = R[i]
    st = sorted(((int(r[j]), h[j]) for j in stalls), reverse=True)[:2]
    print("  %5d %-64s %6s %9s %s" % (i, r[1].strip()[:64], r[si], r[ie], st))

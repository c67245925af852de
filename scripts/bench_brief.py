"""Prints the few numbers of a bench.py JSON line that matter when comparing builds."""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r = d["roofline"]
print("us/step %.2f  frac %.4f  isolated us %.2f  frac_isolated %.4f  e2e %.3g" % (
    d["ms_per_step"] * 1e3, r["frac"], r.get("kernel_ms_isolated_mean", 0) * 1e3, r.get("frac_isolated", 0),
    d["e2e"]["value"]))

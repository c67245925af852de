"""Prints per-tile phase durations from a HLSMAP_TRACE dump (see trace_run.py)."""
import sys
import numpy as np

t = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(-1, 16).astype(np.int64)
print(len(t), "tiles")
t0 = t[:, 0][t[:, 0] > 0].min()


def ph(a, b, name):
    m = (t[:, a] > 0) & (t[:, b] > 0)
    d = (t[m, b] - t[m, a]) / 1e3
    if len(d) == 0:
        return
    print("%-28s med %6.2f mean %6.2f p90 %6.2f us" % (name, np.median(d), d.mean(), np.percentile(d, 90)))


ph(0, 6, "query load")
ph(6, 1, "desc+stage+bounds")
ph(1, 8, "scan")
ph(8, 2, "publish")
ph(0, 2, "search total")
ph(3, 4, "look-back")
ph(4, 5, "offsets+emit")
ph(3, 5, "emit total")
ph(2, 3, "published->emit start")
if (t[:, 13] > 0).any() and (t[:, 14] > 0).any():
    ph(6, 13, "  descriptors")
    ph(13, 14, "  reduce+stage wait")
    ph(14, 1, "  bounds")
if (t[:, 9] > 0).any():
    ph(4, 9, "  offsets+records+bitmap")
    ph(9, 10, "  full rounds (warp 0)")
    ph(10, 11, "  tail round (warp 0)")
    ph(11, 12, "  barrier after the window")
print("kernel span us", (t[:, 5].max() - t0) / 1e3)

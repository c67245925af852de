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
if (t[:, 13] == 0).all() or True:
    ph(3, 14, "  warp 0's walk")
    ph(14, 4, "  then the barrier")
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
d = t[:, 7]
if ((d > 0) & (d < 1 << 20)).any():      # look-back diagnostics of the trace build: windows walked | polls << 8
    w, po = d & 255, d >> 8
    print("look-back windows walked: mean %.2f, share with 1 window %.2f, max %d; polls mean %.2f, share with none %.2f"
          % (w[1:].mean(), (w[1:] == 1).mean(), w.max(), po[1:].mean(), (po[1:] == 0).mean()))
if (t[:, 13] > 0).any() and (t[:, 14] > 0).any() and (t[:, 7] < (1 << 20)).all():    # SMALL kernel: slot 13 = the prefix warp's pass
    ph(2, 13, "published -> prefix warp")
    ph(13, 3, "prefix warp -> emit start")
    m = (t[:, 13] > 0) & (t[:, 3] > 0)
    print("share of tiles whose predecessor was prefixed 1 us before their emit start: %.2f"
          % ((t[1:, 3][m[1:]] - t[:-1, 13][m[1:]]) > 1000).mean())

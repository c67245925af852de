"""Times hlsmap_views_from_fragments_dev and hlsmap_get_next_segment_view_dev on the
config-3 live windows, inputs resident in HBM.  Usage: python scripts/run_fragment_views.py [out.json]"""
import ctypes
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hlsjs_p2p_wrapper_b200 import synth, _native as N           # noqa: E402
from hlsjs_p2p_wrapper_b200.engine import Context                # noqa: E402

ctx, lib, dev = Context(0), N.load(), torch.device("cuda", 0)
tables, _ = synth.config3(n_players=100_000)
m = ctx.upload(tables)
n = 16_000_000
rng = np.random.default_rng(1)
pl = rng.integers(0, 100_000, n).astype(np.int32)
lv = rng.integers(0, 3, n).astype(np.int32)
ix = rng.integers(0, 6, n).astype(np.int32)
want = (tables.sn[(pl.astype(np.int64) * 3 + lv) * 6] + ix).astype(np.int32)
d = [torch.from_numpy(a).to(dev) for a in (pl, lv, ix, want)]
o_sn = torch.empty(n, dtype=torch.int32, device=dev)
o_tm = torch.empty(n, dtype=torch.float64, device=dev)
o_st, o_fd = (torch.empty(n, dtype=torch.uint8, device=dev) for _ in range(2))
stream = torch.cuda.Stream()
v = lambda x: ctypes.c_void_p(x.data_ptr())
sp = ctypes.c_void_p(stream.cuda_stream)


def timed(fn):
    ts = []
    for i in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(stream)
        fn()
        e1.record(stream)
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


ms_a = timed(lambda: N.check(lib.hlsmap_views_from_fragments_dev(m._h, n, v(d[0]), v(d[1]), v(d[2]), v(o_sn), v(o_tm),
                                                                 v(o_st), sp)))
ok_a = bool(torch.equal(o_sn, d[3]) and int(o_st.sum()) == 0)
ms_b = timed(lambda: N.check(lib.hlsmap_get_next_segment_view_dev(m._h, n, v(d[0]), v(d[1]), v(d[3]), v(o_sn), v(o_tm),
                                                                  v(o_fd), v(o_st), sp)))
found = int(o_fd.sum())
out = dict(queries=n, views_from_fragments_ms=ms_a, views_from_fragments_per_s=n / ms_a * 1e3,
           views_from_fragments_GBps=n * 25 / ms_a / 1e6, sn_as_expected=ok_a, next_segment_view_ms=ms_b,
           next_segment_view_per_s=n / ms_b * 1e3, next_found=found, next_expected_found=int((ix < 5).sum()))
print(json.dumps(out))
if len(sys.argv) > 1:
    json.dump(out, open(sys.argv[1], "w"), indent=1)

"""Times viewToString in batch (hlsmap_view_format_dev, columns resident in HBM).  Usage: python scripts/run_format.py"""
import ctypes
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hlsjs_p2p_wrapper_b200 import _native as N                  # noqa: E402
from hlsjs_p2p_wrapper_b200.engine import Context                # noqa: E402

ctx, lib, dev = Context(0), N.load(), torch.device("cuda", 0)
n = 16_000_000
rng = np.random.default_rng(1)
cols = [torch.from_numpy(a).to(dev) for a in (rng.integers(0, 8, n).astype(np.int32), rng.integers(0, 2, n).astype(np.int32),
                                              rng.integers(0, 2 ** 31 - 1, n).astype(np.int32))]
d_off = torch.empty(n + 1, dtype=torch.int32, device=dev)
d_ch = torch.empty(36 * n, dtype=torch.uint8, device=dev)
stream = torch.cuda.Stream()
v = lambda x: ctypes.c_void_p(x.data_ptr())
total = ctypes.c_int64(0)
ts = []
for i in range(7):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    N.check(lib.hlsmap_view_format_dev(ctx._h, n, v(cols[0]), v(cols[1]), v(cols[2]), 1, v(d_off), v(d_ch), 36 * n,
                                       ctypes.c_void_p(stream.cuda_stream), None))
    e1.record(stream)
    torch.cuda.synchronize()
    if i >= 2:
        ts.append(e0.elapsed_time(e1))
N.check(lib.hlsmap_view_format_dev(ctx._h, n, v(cols[0]), v(cols[1]), v(cols[2]), 1, v(d_off), v(d_ch), 36 * n,
                                   ctypes.c_void_p(stream.cuda_stream), ctypes.byref(total)))
ms = float(np.median(ts))
print(json.dumps({"views": n, "chars": total.value, "ms": ms, "views_per_s": n / ms * 1e3,
                  "GBps": (16 * n + total.value) / ms / 1e6}))

#!/usr/bin/env python
"""One 50 000-playlist chunk of BASELINE.json configs[4] through the list form, device-resident,
timed with CUDA events (the general search kernel: tables far beyond shared memory).  Used for
ncu captures and quick timings of that kernel.   usage: run_config5_piece.py [chunk] [reps]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import hlsjs_p2p_wrapper_b200 as pk
from hlsjs_p2p_wrapper_b200 import synth

c = int(sys.argv[1]) if len(sys.argv) > 1 else 0
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
ctx = pk.Context(0)
dev = torch.device("cuda:0")
t, q = synth.config5_chunk(c)
m = ctx.upload(t)
nq = len(q)
d = {k: torch.from_numpy(getattr(q, k)).to(dev) for k in ("playlist", "level", "begin", "duration")}
_, total = m.count_segments(q.playlist, q.level, q.begin, q.duration)
off = torch.empty(nq + 1, dtype=torch.int32, device=dev)
sn = torch.empty(total, dtype=torch.int32, device=dev)
tm = torch.empty(total, dtype=torch.float64, device=dev)
stream = torch.cuda.Stream(device=dev)
ms = []
for i in range(reps + 2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        e0.record()
        m.get_segment_list_dev(nq, d["playlist"].data_ptr(), d["level"].data_ptr(), d["begin"].data_ptr(),
                               d["duration"].data_ptr(), off.data_ptr(), sn.data_ptr(), tm.data_ptr(), total,
                               stream=stream.cuda_stream)
        e1.record()
    torch.cuda.synchronize()
    if i >= 2:
        ms.append(e0.elapsed_time(e1))
alg = 28.0 * nq + 12.0 * total + 12.0 * t.n_fragments
peaks = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
peak = json.load(open(peaks))["hbm_gbs"] if os.path.exists(peaks) else 6650.0
print(json.dumps({"chunk": c, "queries": nq, "hits": int(total), "fragments": t.n_fragments,
                  "ms_median": float(np.median(ms)), "ms_min": float(np.min(ms)),
                  "roofline_frac": alg / (float(np.median(ms)) * 1e-3) / 1e9 / peak}))

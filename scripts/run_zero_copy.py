#!/usr/bin/env python
"""Experiment: the list form writing its results straight into pinned host memory (zero-copy stores
over PCIe) instead of device buffers + D2H copies.  configs[1], 10^6 queries.
usage: run_zero_copy.py [steps]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import hlsjs_p2p_wrapper_b200 as pk
from hlsjs_p2p_wrapper_b200 import synth

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
ctx = pk.Context(0)
dev = torch.device("cuda:0")
t, q = synth.config2(seed=2)
m = ctx.upload(t)
nq = len(q)
_, total = m.count_segments(q.playlist, q.level, q.begin, q.duration)
pins = {k: ctx.pinned(n, dt) for k, n, dt in (("pl", nq, np.int32), ("lv", nq, np.int32), ("bg", nq, np.float64),
                                              ("du", nq, np.float64), ("off", nq + 1, np.int32),
                                              ("sn", total, np.int32), ("tm", total, np.float64))}
pins["pl"].array[:], pins["lv"].array[:], pins["bg"].array[:], pins["du"].array[:] = q.playlist, q.level, q.begin, q.duration
d = {k: torch.empty(nq, dtype=dt, device=dev) for k, dt in (("pl", torch.int32), ("lv", torch.int32),
                                                            ("bg", torch.float64), ("du", torch.float64))}
stream = torch.cuda.Stream(device=dev)
hp = lambda k: pins[k].array.ctypes.data
out = {}


def step_zero_copy(queries_from_host):
    with torch.cuda.stream(stream):
        if not queries_from_host:
            for k in ("pl", "lv", "bg", "du"):
                d[k].copy_(torch.from_numpy(pins[k].array), non_blocking=True)
        src = (lambda k: hp(k)) if queries_from_host else (lambda k: d[k].data_ptr())
        r = m.get_segment_list_dev(nq, src("pl"), src("lv"), src("bg"), src("du"), hp("off"), hp("sn"), hp("tm"),
                                   total, stream=stream.cuda_stream, sync=True)
    return r


for name, qh in (("queries_h2d_results_zero_copy", False), ("all_zero_copy", True)):
    for _ in range(2):
        r = step_zero_copy(qh)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        r = step_zero_copy(qh)
    dt = (time.perf_counter() - t0) / steps
    assert r.total_hits == total and int(pins["off"].array[-1]) == total
    out[name] = {"ms_per_step": dt * 1e3, "queries_per_s": nq / dt, "GBps": (28 * nq + 12 * total) / dt / 1e9}

# the library's own host-pointer path for comparison
for _ in range(2):
    m.get_segment_list_into(pins["pl"].array, pins["lv"].array, pins["bg"].array, pins["du"].array,
                            pins["off"].array, pins["sn"].array, pins["tm"].array)
t0 = time.perf_counter()
for _ in range(steps):
    m.get_segment_list_into(pins["pl"].array, pins["lv"].array, pins["bg"].array, pins["du"].array,
                            pins["off"].array, pins["sn"].array, pins["tm"].array)
dt = (time.perf_counter() - t0) / steps
out["library_pipeline"] = {"ms_per_step": dt * 1e3, "queries_per_s": nq / dt}
import oracle
ref = oracle.get_segment_list_batch(t.playlist_base, t.track_off, t.track_state, t.start, t.sn, q.playlist, q.level, q.begin, q.duration)
step_zero_copy(False)
out["zero_copy_bit_exact"] = bool(np.array_equal(pins["sn"].array, ref[2]) and np.array_equal(pins["tm"].array.view(np.uint64), ref[3].view(np.uint64)))
print(json.dumps(out))

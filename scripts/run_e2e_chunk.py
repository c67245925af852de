import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hlsjs_p2p_wrapper_b200 as pk
from hlsjs_p2p_wrapper_b200 import synth
ctx = pk.Context(0)
t, q = synth.config2(seed=2)
m = ctx.upload(t)
nq = len(q)
_, total = m.count_segments(q.playlist, q.level, q.begin, q.duration)
pins = {k: ctx.pinned(n, dt) for k, n, dt in (("pl", nq, np.int32), ("lv", nq, np.int32), ("bg", nq, np.float64),
        ("du", nq, np.float64), ("off", nq + 1, np.int32), ("sn", total, np.int32), ("tm", total, np.float64))}
pins["pl"].array[:], pins["lv"].array[:], pins["bg"].array[:], pins["du"].array[:] = q.playlist, q.level, q.begin, q.duration
def step():
    return m.get_segment_list_into(pins["pl"].array, pins["lv"].array, pins["bg"].array, pins["du"].array,
                                   pins["off"].array, pins["sn"].array, pins["tm"].array)
for _ in range(3): step()
ts=[]
for _ in range(10):
    t0=time.perf_counter(); r=step(); ts.append(time.perf_counter()-t0)
print(os.environ.get("HLSMAP_HOST_CHUNK","default"), "ms median %.3f min %.3f launches %d" % (1e3*np.median(ts), 1e3*min(ts), r.kernel_launches))

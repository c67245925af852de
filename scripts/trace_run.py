"""Experiment: dump per-tile phase timestamps of one search-kernel launch.  Needs the
diagnostics library: run with HLSMAP_TRACE_BUILD=1 in the environment (builds and loads
csrc/libhlsmap_trace.so; the product library carries no trace code)."""
import os, sys
os.environ["HLSMAP_TRACE_BUILD"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hlsjs_p2p_wrapper_b200 as pk
from hlsjs_p2p_wrapper_b200 import synth
from hlsjs_p2p_wrapper_b200 import build as _b
_b.build()
out = sys.argv[1]
ctx = pk.Context(0)
t, q = synth.config5(n_playlists=20000) if os.environ.get('TRACE_WORKLOAD') == 'config5' else synth.config2()
m = ctx.upload(t)
dev = torch.device("cuda:0")
d = {k: torch.from_numpy(getattr(q, k)).to(dev) for k in ("playlist", "level", "begin", "duration")}
nq = len(q)
_, total = m.count_segments(q.playlist, q.level, q.begin, q.duration)
off = torch.empty(nq + 1, dtype=torch.int32, device=dev)
sn = torch.empty(total, dtype=torch.int32, device=dev)
tm = torch.empty(total, dtype=torch.float64, device=dev)
st = torch.cuda.Stream()
def go():
    m.get_segment_list_dev(nq, d["playlist"].data_ptr(), d["level"].data_ptr(), d["begin"].data_ptr(),
                           d["duration"].data_ptr(), off.data_ptr(), sn.data_ptr(), tm.data_ptr(), total,
                           stream=st.cuda_stream, sync=True)
for _ in range(5):
    go()
os.environ["HLSMAP_TRACE"] = out
go()
del os.environ["HLSMAP_TRACE"]
print("trace written", os.path.getsize(out))

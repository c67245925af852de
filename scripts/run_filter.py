"""Times the filter path (resolve_filter_kernel: the reference's linear loop, used when a table is
unsorted or holds a NaN) on the bench workload, forced with hlsmap_map_set_path, and checks it against
the search path.  Usage: python scripts/run_filter.py [out.json]"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hlsjs_p2p_wrapper_b200 import synth, _native as N           # noqa: E402
from hlsjs_p2p_wrapper_b200.engine import Context                # noqa: E402

ctx, dev = Context(0), torch.device("cuda", 0)
out = []
for name, n_frag, nq in (("config2: 3 levels x 1000 fragments", 1000, 1_000_000), ("3 levels x 100 fragments", 100, 1_000_000)):
    tables, q = synth.config2(n_frag=n_frag, n_queries=nq)
    m = ctx.upload(tables)
    ref = m.get_segment_list(q.playlist, q.level, q.begin, q.duration)
    d = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (q.playlist, q.level, q.begin, q.duration)]
    cap = int(ref.total_hits)
    d_off = torch.empty(nq + 1, dtype=torch.int32, device=dev)
    d_sn = torch.empty(cap, dtype=torch.int32, device=dev)
    d_tm = torch.empty(cap, dtype=torch.float64, device=dev)
    stream = torch.cuda.Stream()
    m.set_path(N.PATH_FILTER)
    ts = []
    for i in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(stream)
        m.get_segment_list_dev(nq, d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(), d_off.data_ptr(),
                               d_sn.data_ptr(), d_tm.data_ptr(), cap, stream=stream.cuda_stream)
        e1.record(stream)
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(e0.elapsed_time(e1))
    same = (np.array_equal(d_off.cpu().numpy(), ref.offsets) and np.array_equal(d_sn.cpu().numpy(), ref.sn)
            and np.array_equal(d_tm.cpu().numpy(), ref.time))
    ms = float(np.median(ts))
    out.append(dict(case=name, queries=nq, hits=cap, ms=ms, queries_per_s=nq / ms * 1e3, equals_search_path=bool(same),
                    fragments_compared_per_s=nq * n_frag / ms * 1e3))
    print(json.dumps(out[-1]))
    m.close()
if len(sys.argv) > 1:
    json.dump(out, open(sys.argv[1], "w"), indent=1)

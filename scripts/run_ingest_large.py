"""Playlist ingest at a size where every index is far beyond one block of the scan: 80 000 VOD
playlists x 800 fragments (2 GB of text, 6.4e7 fragments), through the host-pointer entry, checked
against the oracle.  Usage: python scripts/run_ingest_large.py [n_tracks]"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle                                                    # noqa: E402
from hlsjs_p2p_wrapper_b200 import synth                         # noqa: E402
from hlsjs_p2p_wrapper_b200.engine import Context                # noqa: E402

n_tracks = int(sys.argv[1]) if len(sys.argv) > 1 else 80_000
ctx = Context(0)
text, off = synth.m3u8_block(n_tracks=n_tracks, n_frag=800)
t0 = time.perf_counter()
got = ctx.parse_m3u8(text, off)
t_gpu = time.perf_counter() - t0
t0 = time.perf_counter()
exp = oracle.parse_m3u8(text, off, threads=os.cpu_count() or 1)
t_cpu = time.perf_counter() - t0
ok = all(np.array_equal(got[k], exp[k]) for k in ("track_off", "live", "sn")) and \
    np.array_equal(got["start"].view(np.int64), exp["start"].view(np.int64)) and \
    np.array_equal(got["duration"].view(np.int64), exp["duration"].view(np.int64))
print(json.dumps(dict(text_GB=len(text) / 1e9, tracks=n_tracks, fragments=int(exp["track_off"][-1]), parity=bool(ok),
                      host_call_s=t_gpu, cpu_oracle_s=t_cpu, kernel_launches=int(got["result"].kernel_launches))))

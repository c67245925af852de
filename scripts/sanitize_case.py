#!/usr/bin/env python
"""A small pass over every kernel for compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool memcheck python scripts/sanitize_case.py
Sizes are tiny on purpose (the tools slow kernels down by 10-100x)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle
import hlsjs_p2p_wrapper_b200 as pk
from hlsjs_p2p_wrapper_b200 import synth, _native as N
from hlsjs_p2p_wrapper_b200.tables import FragmentTables, QueryBatch, TRACK_PARSED

ctx = pk.Context(0)


def check(t, q, paths):
    ref = oracle.get_segment_list_batch(t.playlist_base, t.track_off, t.track_state, t.start, t.sn,
                                        q.playlist, q.level, q.begin, q.duration)
    m = ctx.upload(t)
    for path in paths:
        m.set_path(path)
        r = m.get_segment_list(q.playlist, q.level, q.begin, q.duration)
        assert np.array_equal(r.offsets.astype(np.int64), ref[0]) and np.array_equal(r.sn, ref[2])
        assert np.array_equal(r.time.view(np.uint64), ref[3].view(np.uint64))
    f, c, s = m.get_segment_range(q.playlist, q.level, q.begin, q.duration)
    assert np.array_equal(c, np.diff(ref[0]))
    m.get_segment_duration(q.playlist, q.level)
    m.close()


t, q = synth.config2(n_queries=3001)                       # SMALL map: staged tables + hint grid
check(t, q, (N.PATH_SEARCH, N.PATH_FILTER))
t, q = synth.config5(n_playlists=60, queries_per_playlist=40)   # window-staged search
check(t, q, (N.PATH_SEARCH, N.PATH_FILTER))
t, q = synth.config3(n_players=700)
check(t, q, (N.PATH_SEARCH,))
t4, _, _ = synth.config4(n_views=8)                        # 80k fragments: windows too large -> L2 search
rng = np.random.default_rng(0)
q = QueryBatch(np.zeros(1500, np.int32), rng.integers(0, 8, 1500).astype(np.int32),
               rng.integers(-100, 70000, 1500).astype(np.float64), np.where(rng.random(1500) < 0.05, 1e6, 30.0))
check(t4, q, (N.PATH_SEARCH,))
lv, ui, sn = (rng.integers(0, 1 << 31, 4099).astype(np.int32) for _ in range(3))
wire = ctx.view_pack(lv, ui, sn)
assert wire.tobytes() == oracle.view_pack(lv, ui, sn).tobytes()
ctx.view_unpack(wire)
ctx.view_is_equal(lv, ui, sn, lv, ui, sn)
ctx.view_is_in_track(lv, ui, lv, ui)
ctx.view_roundtrip(lv, ui, sn)
o, c = ctx.view_format(lv, ui, sn)
eo, ec = oracle.view_format(lv, ui, sn)
assert c.tobytes() == ec.tobytes()
ctx.close()
print("sanitize case ok")

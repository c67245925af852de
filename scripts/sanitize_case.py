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


def fused(t, gen):
    """device-pointer batches through the batches call: fused into one launch (mode 1) and one launch each (mode 2)"""
    import torch
    dev = torch.device("cuda:0")
    m = ctx.upload(t)
    sets = []
    for i, nq in enumerate((3001, 700, 1, 2049, 5000)):
        qq = gen(i, nq)
        ref = oracle.get_segment_list_batch(t.playlist_base, t.track_off, t.track_state, t.start, t.sn,
                                            qq.playlist, qq.level, qq.begin, qq.duration)
        H = int(ref[0][-1])
        sets.append({"ref": ref, "H": H, "nq": nq,
                     "pl": torch.from_numpy(qq.playlist).to(dev), "lv": torch.from_numpy(qq.level).to(dev),
                     "bg": torch.from_numpy(qq.begin).to(dev), "du": torch.from_numpy(qq.duration).to(dev),
                     "off": torch.empty(nq + 1, dtype=torch.int32, device=dev),
                     "sn": torch.empty(max(H, 1), dtype=torch.int32, device=dev),
                     "tm": torch.empty(max(H, 1), dtype=torch.float64, device=dev)})
    batches = [{"nq": s["nq"], "d_playlist": s["pl"].data_ptr(), "d_level": s["lv"].data_ptr(),
                "d_begin": s["bg"].data_ptr(), "d_duration": s["du"].data_ptr(), "d_offsets": s["off"].data_ptr(),
                "d_sn_out": s["sn"].data_ptr(), "d_time_out": s["tm"].data_ptr(), "capacity": s["H"]} for s in sets]
    for mode in (1, 2):
        ctx.set_batch_overlap(mode)
        for s in sets:
            s["off"].fill_(-1)
        torch.cuda.synchronize()
        assert m.get_segment_list_dev_batches(batches) == len(sets)
        m.last_result()
        torch.cuda.synchronize()
        for s in sets:
            assert np.array_equal(s["off"].cpu().numpy().astype(np.int64), s["ref"][0])
            assert np.array_equal(s["sn"][:s["H"]].cpu().numpy(), s["ref"][2])
            assert np.array_equal(s["tm"][:s["H"]].cpu().numpy().view(np.uint64), s["ref"][3].view(np.uint64))
    ctx.set_batch_overlap(False)
    m.close()


t, _ = synth.config2(n_queries=4)
fused(t, lambda i, nq: synth.config2(seed=500 + i, n_queries=nq)[1])
t, qa = synth.config5(n_playlists=60, queries_per_playlist=120)


def pick(i, nq):
    j = np.sort(np.random.RandomState(i).randint(0, len(qa), size=nq))
    return QueryBatch(qa.playlist[j], qa.level[j], qa.begin[j], qa.duration[j])


fused(t, pick)
ctx.close()
print("sanitize case ok")

"""Sharding by playlist on the GPUs that are really there (media-map.js:28,41: a query
touches one playlist's one level, so playlists shard without any exchange).

`ShardedMap` puts a contiguous playlist range on every visible GPU (one context and one host thread
per GPU), routes a batch by playlist id, and gathers on the host; the result must be bit-for-bit the
CPU oracle's on the unsharded tables.  With one visible GPU the multi-device cases are skipped (two
contexts on one GPU are covered in test_gpu_ingest.py); the 10^8-query corpus of BASELINE.json
configs[4] is checked chunk by chunk on cuda:0 whatever the box."""
import os
import threading

import numpy as np
import pytest

import oracle
import hlsjs_p2p_wrapper_b200 as pk
from hlsjs_p2p_wrapper_b200 import synth, shard, _native as N
from hlsjs_p2p_wrapper_b200.tables import QueryBatch

pytestmark = pytest.mark.gpu


def _oracle(t, q, threads=8):
    return oracle.get_segment_list_batch(t.playlist_base, t.track_off, t.track_state, t.start, t.sn,
                                         q.playlist, q.level, q.begin, q.duration, threads=threads)


def test_sharded_map_over_every_visible_gpu():
    import torch
    n = min(torch.cuda.device_count(), 8)
    if n < 2:
        pytest.skip("one GPU visible")
    t, q = synth.config5(seed=71, n_playlists=6000, queries_per_playlist=50)
    # the batch arrives in arbitrary order, with a few queries on playlists that do not exist
    perm = np.argsort(synth._stream(72, 0).u32(len(q)), kind="stable")
    pl = q.playlist[perm].copy()
    pl[::997] = -1
    pl[5::1499] = t.n_playlists + 3
    q = QueryBatch(pl, q.level[perm], q.begin[perm], q.duration[perm])
    ctxs = [pk.Context(i) for i in range(n)]
    sm = shard.ShardedMap(ctxs, t, q)
    assert len(sm.maps) == n and sm.bounds[0] == 0 and sm.bounds[-1] == t.n_playlists
    assert sorted(c.device for c in ctxs) == list(range(n))
    for _ in range(2):
        off, status, sn, tm = sm.get_segment_list(q)
        o_off, o_st, o_sn, o_tm = _oracle(t, q)
        assert np.array_equal(off, o_off) and np.array_equal(status, o_st)
        assert np.array_equal(sn, o_sn) and np.array_equal(tm.view(np.uint64), o_tm.view(np.uint64))
    sm.close()
    for c in ctxs:
        c.close()


def test_sharded_map_hands_a_workers_error_to_the_caller():
    import torch
    n = max(2, min(torch.cuda.device_count(), 8))
    t, q = synth.config5(seed=73, n_playlists=400, queries_per_playlist=20)
    ctxs = [pk.Context(i % torch.cuda.device_count()) for i in range(n)]
    sm = shard.ShardedMap(ctxs, t, q)
    sm.maps[1].close()                    # the second shard's map is gone: its worker must fail loudly
    with pytest.raises(Exception):
        sm.get_segment_list(q)
    sm.maps[1] = ctxs[1].upload(t.slice_playlists(int(sm.bounds[1]), int(sm.bounds[2])))
    off, status, sn, tm = sm.get_segment_list(q)
    o_off, o_st, o_sn, o_tm = _oracle(t, q)
    assert np.array_equal(off, o_off) and np.array_equal(sn, o_sn)
    sm.close()
    for c in ctxs:
        c.close()


@pytest.mark.parametrize("nq", [1, 3, 4, 31, 255, 1023, 1024, 1025, 4099, 300_001])
def test_range_form_on_a_shared_memory_resident_map(ctx, nq):
    """hlsmap_get_segment_range on a map small enough to be staged whole (range_small_kernel: the
    list form's staging and hint grid, no scan): first index, count and status against the oracle,
    at ragged sizes around its 4-queries-per-thread, 1024-queries-per-tile shape."""
    t, q = synth.config2(seed=50 + nq % 5, n_queries=nq)
    m = ctx.upload(t)
    assert m.info()["stage_all"] == 1
    first, count, status = m.get_segment_range(q.playlist, q.level, q.begin, q.duration)
    rf, rc, rs = oracle.get_segment_range_batch(t.playlist_base, t.track_off, t.track_state, t.start,
                                                q.playlist, q.level, q.begin, q.duration)
    assert np.array_equal(first, rf) and np.array_equal(count, rc) and np.array_equal(status, rs)
    m.close()


def test_range_form_small_map_missing_and_unparsed_levels(ctx):
    t = synth.hls_mock_tables()                       # levels 0, 2 unparsed; level 1 parsed (test/mocks/hls.js)
    m = ctx.upload(t)
    lv = np.array([0, 1, 2, 3, -1, 1, 1], np.int32)
    bg = np.array([0.0, 365.0, 0.0, 0.0, 0.0, float("nan"), -1e9])
    du = np.array([10.0, 33.0, 10.0, 10.0, 10.0, 5.0, float("inf")])
    pl = np.array([0, 0, 0, 0, 0, 0, 7], np.int32)
    first, count, status = m.get_segment_range(pl, lv, bg, du)
    rf, rc, rs = oracle.get_segment_range_batch(t.playlist_base, t.track_off, t.track_state, t.start, pl, lv, bg, du)
    assert np.array_equal(first, rf) and np.array_equal(count, rc) and np.array_equal(status, rs)
    assert list(count[:2]) == [0, 3] and first[1] == 12          # test/media-map.js:46-58: sn 37, 38, 39
    m.close()


def test_config5_full_size_1e8_queries_zero_mismatches(ctx):
    """BASELINE.json configs[4] at full size: 10^8 getSegmentList queries over 10^6 playlists
    (20 chunks of 50 000 playlists, synth.config5_chunk), every chunk resolved on the GPU through the
    C ABI and on the CPU oracle, compared query by query (counts) and hit by hit (sn, fp64 bit
    pattern of time).  north_star: 0 mismatches over 10^8 queries.  Takes about 90 s."""
    import torch
    dev = torch.device("cuda:0")
    n_chunks = int(os.environ.get("HLSMAP_TEST_CONFIG5_CHUNKS", synth.CONFIG5_CHUNKS))
    threads = max(1, len(os.sched_getaffinity(0)))
    box = {}

    def gen(c):
        box[c] = synth.config5_chunk(c)
    th = threading.Thread(target=gen, args=(0,))
    th.start()
    queries = hits = bad_q = bad_h = 0
    for c in range(n_chunks):
        th.join()
        t, q = box.pop(c)
        if c + 1 < n_chunks:
            th = threading.Thread(target=gen, args=(c + 1,))
            th.start()
        m = ctx.upload(t)
        assert m.info()["all_monotone"] == 1
        nq = len(q)
        d = {k: torch.from_numpy(getattr(q, k)).to(dev) for k in ("playlist", "level", "begin", "duration")}
        _, total = m.count_segments(q.playlist, q.level, q.begin, q.duration)
        off = torch.empty(nq + 1, dtype=torch.int32, device=dev)
        st = torch.empty(nq, dtype=torch.uint8, device=dev)
        sn = torch.empty(total, dtype=torch.int32, device=dev)
        tm = torch.empty(total, dtype=torch.float64, device=dev)
        res = m.get_segment_list_dev(nq, d["playlist"].data_ptr(), d["level"].data_ptr(), d["begin"].data_ptr(),
                                     d["duration"].data_ptr(), off.data_ptr(), sn.data_ptr(), tm.data_ptr(),
                                     total, d_status=st.data_ptr(), sync=True)
        assert res.path == N.PATH_SEARCH and res.total_hits == total
        g_off, g_sn, g_tm = off.cpu().numpy().astype(np.int64), sn.cpu().numpy(), tm.cpu().numpy()
        o_off, o_st, o_sn, o_tm = _oracle(t, q, threads)
        dq = int((np.diff(g_off) != np.diff(o_off)).sum()) + int((st.cpu().numpy() != o_st).sum())
        bad_q += dq
        if dq == 0:
            bad_h += int((g_sn != o_sn).sum() + (g_tm.view(np.uint64) != o_tm.view(np.uint64)).sum())
        queries += nq
        hits += int(total)
        m.close()
        del d, off, st, sn, tm
    assert queries == n_chunks * synth.CONFIG5_CHUNK * synth.CONFIG5_QUERIES_PER_PLAYLIST
    assert (bad_q, bad_h) == (0, 0), "%d mismatched queries, %d mismatched hits over %d queries" % (bad_q, bad_h, queries)
    print("config5 full size: %d queries, %d hits, 0 mismatches" % (queries, hits))

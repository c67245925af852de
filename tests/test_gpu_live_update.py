"""Live playlists: hlsmap_map_slide_tracks keeps the HBM copy of the
fragment tables current when windows slide, the way the reference sees a reloaded
hls.levels[..].details.fragments on its next call (media-map.js:28,41).  After every update the
GPU answers must equal the oracle's answers on host tables updated the same way."""
import numpy as np
import pytest

import oracle
from hlsjs_p2p_wrapper_b200 import synth, _native as N
from hlsjs_p2p_wrapper_b200.tables import FragmentTables, QueryBatch

pytestmark = pytest.mark.gpu


def _oracle(t, q):
    return oracle.get_segment_list_batch(t.playlist_base, t.track_off, t.track_state, t.start, t.sn,
                                         q.playlist, q.level, q.begin, q.duration, threads=8)


def _check(m, t_len, host, q):
    """host = dict(start, sn) padded columns with per-track lengths t_len; compare GPU vs oracle."""
    off = np.concatenate([[0], np.cumsum(t_len)]).astype(np.int64)
    start = np.concatenate([host["start"][i][:n] for i, n in enumerate(t_len)])
    sn = np.concatenate([host["sn"][i][:n] for i, n in enumerate(t_len)]).astype(np.int32)
    t = FragmentTables(host["pbase"], off, host["state"], host["urlc"], start, sn, np.ones(len(start)))
    ref = _oracle(t, q)
    res = m.get_segment_list(q.playlist, q.level, q.begin, q.duration)
    assert np.array_equal(res.offsets.astype(np.int64), ref[0])
    assert np.array_equal(res.sn, ref[2])
    assert np.array_equal(res.time.view(np.uint64), ref[3].view(np.uint64))
    return res


@pytest.mark.parametrize("n_players", [7, 3000])          # shared-memory-resident map / windowed map
def test_sliding_windows_follow_the_playlists(ctx, n_players):
    t, q = synth.config3(seed=11, n_players=n_players)
    nt = t.n_tracks
    host = {"pbase": t.playlist_base, "state": t.track_state, "urlc": t.url_count,
            "start": [t.start[t.track_off[i]:t.track_off[i + 1]].copy() for i in range(nt)],
            "sn": [t.sn[t.track_off[i]:t.track_off[i + 1]].copy() for i in range(nt)]}
    lens = np.diff(t.track_off).astype(np.int64)
    m = ctx.upload(t)
    assert m.info()["stage_all"] == (1 if n_players == 7 else 0)
    _check(m, lens, host, q)
    rng = np.random.default_rng(n_players)
    for step in range(3):
        # slide a random subset of tracks by 1..3 fragments (evict front, append at the back)
        pick = np.nonzero(rng.random(nt) < 0.6)[0]
        k = rng.integers(1, 4, len(pick))
        pl, lv = (pick // 3).astype(np.int32), (pick % 3).astype(np.int32)
        aoff = np.concatenate([[0], np.cumsum(k)]).astype(np.int64)
        ns, nsn = np.empty(aoff[-1]), np.empty(aoff[-1], np.int32)
        for j, tr in enumerate(pick):
            last_s, last_sn = host["start"][tr][lens[tr] - 1], host["sn"][tr][lens[tr] - 1]
            add_s = last_s + 2.0 * np.arange(1, k[j] + 1) + rng.integers(0, 512, k[j]) / 1024.0
            add_sn = last_sn + np.arange(1, k[j] + 1)
            ns[aoff[j]:aoff[j + 1]], nsn[aoff[j]:aoff[j + 1]] = add_s, add_sn
            host["start"][tr] = np.concatenate([host["start"][tr][k[j]:lens[tr]], add_s])
            host["sn"][tr] = np.concatenate([host["sn"][tr][k[j]:lens[tr]], add_sn]).astype(np.int32)
        m.slide_tracks(pl, lv, k.astype(np.int32), aoff, ns, nsn, new_dur0=np.full(len(pick), 2.5 + step))
        # queries that follow the windows
        first = np.array([host["start"][int(p) * 3 + int(l)][0] for p, l in zip(q.playlist, q.level)])
        q2 = QueryBatch(q.playlist, q.level, first + rng.integers(-8, 8, len(first)) / 4.0, q.duration)
        res = _check(m, lens, host, q2)
        assert res.path == N.PATH_SEARCH and res.total_hits > 0
        d, st = m.get_segment_duration(pl, lv)
        assert (d == 2.5 + step).all() and (st == 0).all()
    m.close()


def test_shrink_grow_and_unsorted_updates(ctx):
    t, q = synth.config3(seed=12, n_players=5)
    host = {"pbase": t.playlist_base, "state": t.track_state, "urlc": t.url_count,
            "start": [t.start[t.track_off[i]:t.track_off[i + 1]].copy() for i in range(15)],
            "sn": [t.sn[t.track_off[i]:t.track_off[i + 1]].copy() for i in range(15)]}
    lens = np.diff(t.track_off).astype(np.int64)
    m = ctx.upload(t)
    # evicting at the front changes the first fragment: its duration has to come along (the reference
    # would read the new fragments[0].duration, media-map.js:81-87), a missing one is refused
    with pytest.raises(N.HlsMapError) as e:
        m.slide_tracks([1, 2], [0, 1], [4, 6], [0, 0, 0], np.zeros(0), np.zeros(0, np.int32))
    assert e.value.code == N.ERR_INVALID and "new_dur0" in e.value.message
    # shrink track (1,0) to 2 fragments, empty track (2,1), leave the others
    m.slide_tracks([1, 2], [0, 1], [4, 6], [0, 0, 0], np.zeros(0), np.zeros(0, np.int32), new_dur0=[7.5, np.nan])
    d, st = m.get_segment_duration([1, 2], [0, 1])
    assert d[0] == 7.5 and st[0] == N.ST_OK and st[1] == N.ST_EMPTY
    host["start"][3], host["sn"][3] = host["start"][3][4:], host["sn"][3][4:]
    host["start"][7], host["sn"][7] = host["start"][7][:0], host["sn"][7][:0]
    lens[3], lens[7] = 2, 0
    _check(m, lens, host, q)
    # grow back up to the reserved six fragments, not beyond
    s0 = host["start"][3][-1]
    m.slide_tracks([1], [0], [0], [0, 4], s0 + np.arange(1, 5), np.arange(900, 904, dtype=np.int32))
    host["start"][3] = np.concatenate([host["start"][3], s0 + np.arange(1, 5)])
    host["sn"][3] = np.concatenate([host["sn"][3], np.arange(900, 904)]).astype(np.int32)
    lens[3] = 6
    _check(m, lens, host, q)
    with pytest.raises(N.HlsMapError) as e:
        m.slide_tracks([1], [0], [0], [0, 1], [1e9], [1])
    assert e.value.code == N.ERR_CAPACITY
    with pytest.raises(N.HlsMapError):
        m.slide_tracks([1, 1], [0, 0], [1, 1], [0, 0, 0], np.zeros(0), np.zeros(0, np.int32), new_dur0=[1.0, 1.0])   # twice
    # an update that breaks the order: the map must fall back to the filter kernel, still exact
    bad = host["start"][0][0] - 50.0
    m.slide_tracks([0], [0], [1], [0, 1], [bad], [7], new_dur0=[2.0])
    host["start"][0] = np.concatenate([host["start"][0][1:], [bad]])
    host["sn"][0] = np.concatenate([host["sn"][0][1:], [7]]).astype(np.int32)
    assert m.info()["all_monotone"] == 0 and m.monotone_flags()[0] == 0
    q3 = QueryBatch(np.zeros(64, np.int32), np.zeros(64, np.int32),
                    np.linspace(bad - 5, host["start"][0][-2] + 5, 64), np.full(64, 30.0))
    res = m.get_segment_list(q3.playlist, q3.level, q3.begin, q3.duration)
    assert res.path == N.PATH_FILTER
    _check(m, lens, host, q3)
    m.close()


def test_sequence_numbers_that_stop_being_contiguous(ctx):
    """The windowed search kernel computes a hit's sn from its query's first hit only while every
    track numbers its fragments contiguously (proven on the GPU at upload and after every slide).
    A discontinuity -- at upload, or appended later -- must send it back to reading sn[]: answers
    stay bit-for-bit the oracle's, including sn values that wrap around INT32."""
    t, q = synth.config5(seed=21, n_playlists=900, queries_per_playlist=40)     # beyond shared memory
    nt = t.n_tracks
    lens = np.diff(t.track_off).astype(np.int64)
    sn = t.sn.copy()
    # contiguous runs that wrap: first track starts just below INT32_MAX
    o0, n0 = int(t.track_off[0]), int(lens[0])
    sn[o0:o0 + n0] = ((np.arange(n0, dtype=np.int64) + (2 ** 31 - 3) + 2 ** 31) % 2 ** 32 - 2 ** 31).astype(np.int32)
    tw = FragmentTables(t.playlist_base, t.track_off, t.track_state, t.url_count, t.start, sn, t.duration)
    m = ctx.upload(tw)
    assert m.info()["stage_all"] == 0 and m.info()["all_monotone"] == 1
    ref = _oracle(tw, q)
    res = m.get_segment_list(q.playlist, q.level, q.begin, q.duration)
    assert np.array_equal(res.offsets.astype(np.int64), ref[0]) and np.array_equal(res.sn, ref[2])
    assert np.array_equal(res.time.view(np.uint64), ref[3].view(np.uint64))
    # a slide that appends a gap in the numbering (an HLS discontinuity in sn) to a few tracks
    host = {"pbase": t.playlist_base, "state": t.track_state, "urlc": t.url_count,
            "start": [tw.start[t.track_off[i]:t.track_off[i + 1]].copy() for i in range(nt)],
            "sn": [tw.sn[t.track_off[i]:t.track_off[i + 1]].copy() for i in range(nt)]}
    pick = np.array([3, 4, 500, nt - 1])
    k = np.array([2, 1, 3, 2])
    aoff = np.concatenate([[0], np.cumsum(k)]).astype(np.int64)
    ns, nsn = np.empty(aoff[-1]), np.empty(aoff[-1], np.int32)
    for j, tr in enumerate(pick):
        last_s, last_sn = host["start"][tr][lens[tr] - 1], int(host["sn"][tr][lens[tr] - 1])
        add_s = last_s + 3.0 * np.arange(1, k[j] + 1)
        add_sn = last_sn + 7 + np.arange(1, k[j] + 1)                  # the gap
        ns[aoff[j]:aoff[j + 1]], nsn[aoff[j]:aoff[j + 1]] = add_s, add_sn
        host["start"][tr] = np.concatenate([host["start"][tr][k[j]:lens[tr]], add_s])
        host["sn"][tr] = np.concatenate([host["sn"][tr][k[j]:lens[tr]], add_sn]).astype(np.int32)
    m.slide_tracks((pick // 3).astype(np.int32), (pick % 3).astype(np.int32), k.astype(np.int32), aoff, ns, nsn,
                   new_dur0=np.full(len(pick), 3.0))
    _check(m, lens, host, q)
    m.close()
    # and a map that is not contiguous from the start
    sn2 = t.sn.copy()
    sn2[::5] += 1000
    t2 = FragmentTables(t.playlist_base, t.track_off, t.track_state, t.url_count, t.start, sn2, t.duration)
    m = ctx.upload(t2)
    ref = _oracle(t2, q)
    res = m.get_segment_list(q.playlist, q.level, q.begin, q.duration)
    assert np.array_equal(res.offsets.astype(np.int64), ref[0]) and np.array_equal(res.sn, ref[2])
    m.close()

"""View construction from fragments (the loader's `new SegmentView({sn: frag.sn, trackView, time:
frag.start})`, p2p-loader-generator.js:145-147) and the getNextSegmentView EXTENSION (absent from
this reference, CHANGELOG.md:89: parity only against the library's own restatement)."""
import numpy as np
import pytest

import oracle
from oracle import pyref
from hlsjs_p2p_wrapper_b200 import synth, _native as N
from hlsjs_p2p_wrapper_b200.tables import FragmentTables

pytestmark = pytest.mark.gpu


def _tables():
    lv = lambda fr: {"url": ["a", "b"], "details": {"fragments": fr}}
    f = lambda n, sn0, s0=0.0: [{"sn": sn0 + i, "start": s0 + 2.5 * i} for i in range(n)]
    odd = [{"sn": s, "start": 1.0 * i} for i, s in enumerate([7, 3, 3, 9, 7, 1])]     # repeats, unordered
    return FragmentTables.from_hls_levels([[lv(f(40, 1000)), {"url": ["a"]}, lv([]), lv(odd)],
                                           None, [lv(f(3, 5, 9.0))]])


def test_views_from_fragments_match_loader_construction(ctx):
    t = _tables()
    m = ctx.upload(t)
    rng = np.random.default_rng(1)
    n = 4000
    pl = rng.integers(-1, 4, n).astype(np.int32)
    lv = rng.integers(-1, 5, n).astype(np.int32)
    ix = rng.integers(-2, 45, n).astype(np.int32)
    sn, tm, st = m.views_from_fragments(pl, lv, ix)
    esn, etm, est = oracle.views_from_fragments(t.playlist_base, t.track_off, t.track_state, t.start, t.sn, pl, lv, ix)
    assert np.array_equal(st, est) and np.array_equal(sn, esn) and np.array_equal(tm, etm, equal_nan=True)
    assert set(np.unique(st)) == {N.ST_OK, N.ST_UNPARSED, N.ST_MISSING, N.ST_NO_FRAGMENT}
    # the reference's own construction on one fragment
    mock = pyref.HlsMock(3, False, 1)
    tm1 = ctx.upload(synth.hls_mock_tables(3, 1, False))
    s1, t1, st1 = tm1.views_from_fragments([0], [1], [12])
    frag = mock.levels[1]["details"]["fragments"][12]
    sv = pyref.SegmentView({"sn": frag["sn"], "trackView": {"level": 1, "urlId": 1}, "time": frag["start"]})
    assert (int(s1[0]), float(t1[0]), int(st1[0])) == (sv.sn, sv.time, 0)
    tm1.close()
    m.close()


def test_next_segment_view_extension(ctx):
    t = _tables()
    m = ctx.upload(t)
    rng = np.random.default_rng(2)
    n = 6000
    pl = rng.integers(-1, 4, n).astype(np.int32)
    lv = rng.integers(-1, 5, n).astype(np.int32)
    want = np.where(rng.random(n) < 0.5, rng.integers(995, 1045, n), rng.integers(0, 12, n)).astype(np.int32)
    got = m.get_next_segment_view(pl, lv, want)
    exp = oracle.get_next_segment_view(t.playlist_base, t.track_off, t.track_state, t.start, t.sn, pl, lv, want)
    for g, e in zip(got, exp):
        assert np.array_equal(g, e, equal_nan=True)
    assert got[2].any() and not got[2].all()
    # spelled out: contiguous numbering, the last fragment has no successor, repeated sn -> first match
    s, tt, f, _ = m.get_next_segment_view([0, 0, 0, 0, 0], [0, 0, 3, 3, 3], [1000, 1039, 3, 7, 1])
    assert list(f) == [1, 0, 1, 1, 0]
    assert (int(s[0]), float(tt[0])) == (1001, 2.5) and (int(s[2]), float(tt[2])) == (3, 2.0)
    assert (int(s[3]), float(tt[3])) == (3, 1.0)
    m.close()


def test_config3_live_windows_next_view(ctx):
    """BASELINE.json configs[2] as literally worded: getNextSegmentView + getSegmentList."""
    t, q = synth.config3(n_players=5000)
    m = ctx.upload(t)
    first = t.sn[(q.playlist.astype(np.int64) * 3 + q.level) * 6]
    want = (first + np.arange(len(q)) % 7).astype(np.int32)          # the 7th is past the window
    got = m.get_next_segment_view(q.playlist, q.level, want)
    exp = oracle.get_next_segment_view(t.playlist_base, t.track_off, t.track_state, t.start, t.sn,
                                       q.playlist, q.level, want)
    for g, e in zip(got, exp):
        assert np.array_equal(g, e, equal_nan=True)
    assert got[2].sum() == (np.arange(len(q)) % 7 < 5).sum()
    m.close()

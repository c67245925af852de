"""Property tests on the oracle (CPU only): the C restatement agrees with the object-level
Python restatement on arbitrary tables (unsorted, NaN, +-Inf, -0, duplicates), and the
search formulation the CUDA kernel uses equals the reference's filter exactly when the
start column is monotone and NaN-free (hard parts 1-2)."""
import math

import numpy as np
from hypothesis import given, settings, strategies as st

import oracle
from oracle import pyref

specials = [0.0, -0.0, float("inf"), float("-inf"), float("nan"), 0.1, 0.2, 0.3, 0.1 + 0.2,
            1e308, -1e308, 5e-324, 10.0, 20.0, 30.0]
f64 = st.one_of(st.sampled_from(specials), st.floats(allow_nan=True, allow_infinity=True, width=64),
                st.integers(-50, 50).map(float), st.integers(-5000, 5000).map(lambda v: v / 1024.0))


def _pyref_list(start, sn, begin, dur):
    frags = [{"sn": int(s), "start": float(x)} for s, x in zip(sn, start)]
    hls = type("H", (), {"levels": [{"url": ["a"], "details": {"fragments": frags}}]})()
    tv = pyref.TrackView({"level": 0, "urlId": 0})
    got = pyref.MediaMap(hls).getSegmentList(tv, begin, dur)
    return [g.sn for g in got], [g.time for g in got]


def _bits(xs):
    return np.asarray(xs, np.float64).view(np.uint64).tolist()


@settings(max_examples=300, deadline=None)
@given(st.lists(f64, max_size=40), f64, f64)
def test_c_oracle_equals_pyref_on_any_table(start, begin, dur):
    sn = list(range(100, 100 + len(start)))
    esn, etm = _pyref_list(start, sn, begin, dur)
    gsn, gtm = oracle.filter_one(start, sn, begin, dur)
    assert list(gsn) == esn
    assert _bits(gtm) == _bits(etm)          # time is fragment.start bit for bit


def search_formulation(start, begin, dur):
    """The CUDA search kernel's bounds, statement for statement (resolve.cu)."""
    n = len(start)
    ub = begin + dur
    lo, m = 0, n
    while m > 0:
        half = m >> 1
        keep_right = not (begin <= start[lo + half])
        lo, m = (lo + half + 1, m - half - 1) if keep_right else (lo, half)
    a, b, step = lo, n, 1
    while True:
        pr = a + step - 1
        if pr >= n:
            break
        if start[pr] <= ub:
            a, step = pr + 1, step << 1
        else:
            b = pr
            break
    while a < b:
        mid = (a + b) >> 1
        if start[mid] <= ub:
            a = mid + 1
        else:
            b = mid
    return lo, a - lo


sortable = st.one_of(st.sampled_from([0.0, -0.0, float("inf"), float("-inf"), 0.3, 0.1 + 0.2, 1e308]),
                     st.floats(allow_nan=False, allow_infinity=True, width=64),
                     st.integers(-30, 30).map(float))


@settings(max_examples=500, deadline=None)
@given(st.lists(sortable, max_size=60), f64, f64)
def test_search_equals_filter_on_monotone_tables(start, begin, dur):
    start = sorted(start)                      # -0.0 and 0.0 may land in either order: both legal
    sn = list(range(len(start)))
    gsn, gtm = oracle.filter_one(start, sn, begin, dur)
    lo, cnt = search_formulation(start, begin, dur)
    assert cnt == len(gsn)
    if cnt:
        assert list(gsn) == list(range(lo, lo + cnt))
        assert _bits(gtm) == _bits(start[lo:lo + cnt])


@settings(max_examples=200, deadline=None)
@given(st.lists(f64, max_size=30))
def test_is_monotone_matches_definition(start):
    expect = all(not math.isnan(x) for x in start) and \
        all(start[i] >= start[i - 1] for i in range(1, len(start)))
    assert oracle.is_monotone(start) == expect


def test_rounded_upper_bound_is_one_fp64_add():
    """0.1 + 0.2 rounds to 0.30000000000000004: a fragment at exactly that start is kept,
    one at 0.3 too, one at nextafter(sum) is not (media-map.js:44)."""
    s = 0.1 + 0.2
    start = [0.3, s, math.nextafter(s, 1.0)]
    gsn, _ = oracle.filter_one(start, [0, 1, 2], 0.1, 0.2)
    assert list(gsn) == [0, 1]
    assert search_formulation(start, 0.1, 0.2) == (0, 2)


def test_batch_statuses_and_offsets():
    from hlsjs_p2p_wrapper_b200.synth import hls_mock_tables
    t = hls_mock_tables(3, 1, False)
    off, status, sn, tm = oracle.get_segment_list_batch(
        t.playlist_base, t.track_off, t.track_state, t.start, t.sn,
        [0, 0, 0, 0, 1], [1, 0, 4, -1, 1], [365, 365, 365, 365, 365], [33, 33, 33, 33, 33], threads=2)
    assert list(status) == [0, 1, 2, 2, 2]
    assert list(off) == [0, 3, 3, 3, 3, 3]
    assert list(sn) == [37, 38, 39]
    out, st_ = oracle.get_segment_duration_batch(t.playlist_base, t.track_off, t.track_state, t.dur0(),
                                                 [0, 0, 0], [1, 0, 7])
    assert list(st_) == [0, 1, 2] and math.isnan(out[0])

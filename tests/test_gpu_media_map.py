"""The reference's mocha suite for the mapping layer (test/media-map.js,
test/segment-view.js, test/track-view.js), replayed against the drop-in classes whose
getSegmentList / getSegmentDuration run on the GPU through the C ABI.  Expected values
come from tests/golden/reference_vectors.json; the fixture is the restated HlsMock."""
import json
import warnings

import numpy as np
import pytest

from oracle import pyref
from hlsjs_p2p_wrapper_b200 import MediaMap, SegmentView, TrackView, JSError, JSTypeError, undefined

pytestmark = pytest.mark.gpu


def _mock(args):
    n, live, defined = args
    m = pyref.HlsMock(n, live, defined)
    lv = m.levels
    return type("Hls", (), {"levels": None if lv is pyref.UNDEFINED else lv})()


def _expected(v):
    if "expect_sn_range" in v:
        sn = list(range(*v["expect_sn_range"]))
        return sn, [s * v["expect_time_per_sn"] for s in sn]
    return v["expect_sn"], v["expect_time"]


@pytest.mark.parametrize("i", range(4))
def test_get_segment_time(golden, ctx, i):
    v = golden["getSegmentTime"][i]
    mm = MediaMap(_mock(v["mock"]), ctx)
    obj = {"sn": v["view"]["sn"], "trackView": {"level": v["view"]["level"], "urlId": v["view"]["urlId"]}}
    if "time" in v["view"]:
        obj["time"] = v["view"]["time"]
    sv = SegmentView(obj)
    if "throws" in v:
        with pytest.raises(JSError) as e:
            mm.getSegmentTime(sv)
        assert str(e.value) == v["throws"]
    else:
        assert mm.getSegmentTime(sv) == v["expect"]


@pytest.mark.parametrize("i", range(7))
def test_get_segment_list(golden, ctx, i):
    v = golden["getSegmentList"][i]
    mm = MediaMap(_mock(v["mock"]), ctx)
    tv = TrackView(v["track"])
    if "throws" in v:
        with pytest.raises(JSError) as e:
            mm.getSegmentList(tv, v["begin"], v["duration"])
        assert str(e.value) == v["throws"]
        return
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        got = mm.getSegmentList(tv, v["begin"], v["duration"])
    sn, tm = _expected(v)
    expect = [SegmentView({"sn": s, "trackView": tv, "time": float(t)}) for s, t in zip(sn, tm)]
    assert got == expect                                   # should.eql on real instances
    assert all(isinstance(g, SegmentView) and isinstance(g.trackView, TrackView) for g in got)
    assert all(g.trackView is not tv for g in got)         # fresh TrackView copies
    assert [str(x.message) for x in w] == ([v["warns"]] if "warns" in v else [])
    # both kernels give the same answer
    for path in (2, 1):
        mm._device_map().set_path(path)
        assert mm.getSegmentList(tv, v["begin"], v["duration"]) == expect if "warns" not in v else True
    mm._device_map().set_path(0)


def test_get_segment_list_levels_undefined(ctx):
    mm = MediaMap(_mock([0, False, 0]), ctx)
    with pytest.raises(JSTypeError):
        mm.getSegmentList(TrackView({"level": 1, "urlId": 1}), 0, 10)


def test_get_track_list(golden, ctx):
    for v in golden["getTrackList"]:
        got = MediaMap(_mock(v["mock"]), ctx).getTrackList()
        assert len(got) == v["expect_length"]
        if "expect" in v:
            assert [[t.level, t.urlId] for t in got] == v["expect"]
            assert all(isinstance(t, TrackView) for t in got)


def test_get_segment_duration(ctx):
    """media-map.js:81-87 -- no reference test; checked against the restatement."""
    frags = [{"sn": 7, "start": 70.0, "duration": 9.5}, {"sn": 8, "start": 79.5, "duration": 4.0}]
    hls = type("Hls", (), {"levels": [{"url": ["u"]}, {"url": ["u"], "details": {"fragments": frags}},
                                      {"url": ["u"], "details": {"fragments": []}}]})()
    mm, ref = MediaMap(hls, ctx), pyref.MediaMap(hls)
    sv = SegmentView({"sn": 8, "trackView": {"level": 1, "urlId": 0}, "time": 79.5})
    assert mm.getSegmentDuration(sv) == ref.getSegmentDuration(sv) == 9.5     # first fragment's
    empty = SegmentView({"sn": 8, "trackView": {"level": 2, "urlId": 0}})
    with pytest.raises(JSError) as e:
        mm.getSegmentDuration(empty)
    assert str(e.value) == "All segments should have a duration"
    with pytest.raises(pyref.JSError):
        ref.getSegmentDuration(empty)
    with pytest.raises(JSTypeError):
        mm.getSegmentDuration(SegmentView({"sn": 8, "trackView": {"level": 0, "urlId": 0}}))
    # the mocha fixture has no `duration` field: the reference returns undefined
    mock = _mock([3, False, 1])
    assert MediaMap(mock, ctx).getSegmentDuration(
        SegmentView({"sn": 30, "trackView": {"level": 1, "urlId": 1}})) is undefined


def test_live_reload_is_seen(ctx):
    """The reference reads hls.levels on every call (media-map.js:28,41)."""
    hls = _mock([3, False, 0])
    mm = MediaMap(hls, ctx)
    tv = TrackView({"level": 1, "urlId": 1})
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        assert mm.getSegmentList(tv, 365, 33) == []
    hls.levels[1]["details"] = {"live": False, "fragments": [{"sn": s, "start": s * 10} for s in range(25, 200)]}
    assert [g.sn for g in mm.getSegmentList(tv, 365, 33)] == [37, 38, 39]


def test_segment_view_vectors(golden):
    g = golden["segmentView"]
    for v in g["json_roundtrip"]:
        sv = SegmentView(v["ctor"])
        s = json.dumps(sv.toJSON(), separators=(",", ":"))
        assert s == v["derived_json"]
        assert SegmentView(json.loads(s)).isEqual(sv) is v["expect_equal"]
    for v in g["arraybuffer"]:
        sv = SegmentView(v["ctor"])
        buf = sv.toArrayBuffer()
        if "expect_byte_length" in v:
            assert isinstance(buf, bytes) and len(buf) == 12 and buf.hex() == v["expect_bytes_hex"]
        if v.get("roundtrip_equal"):
            assert SegmentView.fromArrayBuffer(buf).isEqual(sv)
    for v in g["isInTrack"]:
        assert SegmentView(v["view"]).isInTrack(TrackView(v["track"])) is v["expect"]
    for v in g["isEqual"]:
        assert SegmentView(v["a"]).isEqual(SegmentView(v["b"])) is v["expect"]


def test_track_view_vectors(golden):
    g = golden["trackView"]
    for v in g["isEqual"]:
        assert TrackView(v["a"]).isEqual(TrackView(v["b"])) is v["expect"]
    for v in g["viewToString"]:
        assert (TrackView(v["a"]).viewToString() == TrackView(v["b"]).viewToString()) is v["expect_same"]
    assert TrackView({"level": 0, "urlId": 1}).type == "video"


def test_batch_entry_point(ctx):
    mm = MediaMap(_mock([3, False, 1]), ctx)
    res = mm.getSegmentListBatch([1, 1, 1], [365.0, 10.0, 2100.0], [33.0, 275.0, 3000.0])
    assert list(res.offsets) == [0, 3, 7, 7]
    assert list(res.sn) == [37, 38, 39, 25, 26, 27, 28]
    assert list(res.time) == [370.0, 380.0, 390.0, 250.0, 260.0, 270.0, 280.0]
    with pytest.raises(JSError) as e:
        mm.getSegmentListBatch([1, 4], [365.0, 365.0], [33.0, 33.0])
    assert str(e.value) == "getSegmentList: level doesn't exist"
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        res = mm.getSegmentListBatch([0, 1], [365.0, 365.0], [33.0, 33.0])
    assert [str(x.message) for x in w] == ["getSegmentList: level not parsed yet"]
    assert list(res.status) == [1, 0] and list(res.offsets) == [0, 0, 3]
    res = mm.getSegmentListBatch([1, 4], [365.0, 365.0], [33.0, 33.0], strict=False)
    assert list(res.status) == [0, 2] and res.n_missing == 1


# ---- hls.levels is read live (media-map.js:28,41): the device copy must never answer from stale tables
def _levels(n=40, first_sn=100, dur=4.0):
    frag = lambda i: {"sn": first_sn + i, "start": dur * i, "duration": dur}
    return [{"url": ["a", "b"], "details": {"fragments": [frag(i) for i in range(n)]}},
            {"url": ["a"], "details": {"fragments": [frag(i) for i in range(n)]}}]


def _sns(mm, level, begin, duration):
    return [(v.sn, v.time) for v in mm.getSegmentList(TrackView({"level": level, "urlId": 0}), begin, duration)]


def test_interior_start_nudge_is_seen(ctx):
    """hls.js moves fragment starts in place when it aligns PTS: nothing about the list's length,
    first or last entry changes, and the reference would still see the new value on its next call."""
    hls = type("Hls", (), {"levels": _levels()})()
    mm = MediaMap(hls, ctx)
    assert _sns(mm, 0, 39.0, 6.0) == [(110, 40.0), (111, 44.0)]
    hls.levels[0]["details"]["fragments"][10]["start"] = 38.5          # interior nudge, same object graph
    assert _sns(mm, 0, 39.0, 6.0) == [(111, 44.0)]
    assert _sns(mm, 0, 38.0, 6.5) == [(110, 38.5), (111, 44.0)]
    assert _sns(mm, 1, 39.0, 6.0) == [(110, 40.0), (111, 44.0)]        # the other level is untouched
    # url list and first duration are part of what is compared too
    assert [(t.level, t.urlId) for t in mm.getTrackList()] == [(0, 0), (0, 1), (1, 0)]
    hls.levels[1]["url"].append("c")
    assert [(t.level, t.urlId) for t in mm.getTrackList()] == [(0, 0), (0, 1), (1, 0), (1, 1)]
    hls.levels[0]["details"]["fragments"][0]["duration"] = 3.25
    assert mm.getSegmentDuration(SegmentView({"sn": 1, "trackView": {"level": 0, "urlId": 0}, "time": 0})) == 3.25


def test_sliding_window_is_slid_on_the_gpu_not_reuploaded(ctx):
    hls = type("Hls", (), {"levels": _levels(n=12)})()
    mm = MediaMap(hls, ctx)
    assert _sns(mm, 0, 0.0, 8.0) == [(100, 0.0), (101, 4.0), (102, 8.0)]
    uploads = mm.refreshes["upload"]
    for step in range(1, 6):                     # a live reload: two out at the front, two in at the back
        for lv in hls.levels:
            fr = lv["details"]["fragments"]
            last = fr[-1]
            del fr[:2]
            fr.extend({"sn": last["sn"] + j, "start": last["start"] + 4.0 * j + 0.125 * step, "duration": 4.0 + step}
                      for j in (1, 2))
        got = _sns(mm, 0, 0.0, 1e9)
        want = [(f["sn"], f["start"]) for f in hls.levels[0]["details"]["fragments"]]
        assert got == want
        d = mm.getSegmentDuration(SegmentView({"sn": 1, "trackView": {"level": 1, "urlId": 0}, "time": 0}))
        assert d == hls.levels[1]["details"]["fragments"][0]["duration"]
    assert mm.refreshes["upload"] == uploads and mm.refreshes["slide"] == 5
    # a window that grows beyond what was reserved at upload cannot slide: re-upload, same answers
    fr = hls.levels[0]["details"]["fragments"]
    fr.extend({"sn": fr[-1]["sn"] + j, "start": fr[-1]["start"] + 4.0 * j, "duration": 4.0} for j in (1, 2, 3))
    assert _sns(mm, 0, 0.0, 1e9) == [(f["sn"], f["start"]) for f in fr]
    assert mm.refreshes["upload"] == uploads + 1


def test_get_track_list_throws_on_a_falsy_level(ctx):
    """media-map.js:66-68 reads level.url.length of every entry of hls.levels: a hole throws a
    TypeError there, it does not yield an empty track list for that level."""
    hls = type("Hls", (), {"levels": [_levels()[0], None, _levels()[1]]})()
    mm = MediaMap(hls, ctx)
    with pytest.raises(JSTypeError):
        mm.getTrackList()
    hls.levels[1] = {"url": []}
    assert [(t.level, t.urlId) for t in mm.getTrackList()] == [(0, 0), (0, 1), (2, 0)]

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

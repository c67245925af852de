"""CPU-only: the host-side mirror classes (hlsjs-p2p-wrapper_b200/mapping.py) against the
reference's test/segment-view.js and test/track-view.js vectors and against the object-level
restatement (oracle/pyref.py).  These methods never touch the GPU (they are JS object
semantics, kept in the host language like the reference keeps them in JS)."""
import json

import pytest
from hypothesis import given, settings, strategies as st

from oracle import pyref
from hlsjs_p2p_wrapper_b200 import SegmentView, TrackView, MediaMap, JSError, JSTypeError, undefined


def test_segment_view_vectors(golden):
    g = golden["segmentView"]
    for v in g["json_roundtrip"]:
        sv = SegmentView(v["ctor"])
        s = json.dumps(sv.toJSON(), separators=(",", ":"))
        assert s == v["derived_json"] == pyref.SegmentView(v["ctor"]).json_stringify()
        assert SegmentView(json.loads(s)).isEqual(sv) is v["expect_equal"]
    for v in g["arraybuffer"]:
        sv = SegmentView(v["ctor"])
        buf = sv.toArrayBuffer()
        assert buf == pyref.SegmentView(v["ctor"]).toArrayBuffer()
        if "expect_byte_length" in v:
            assert len(buf) == v["expect_byte_length"] and buf.hex() == v["expect_bytes_hex"]
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
        a, b = TrackView(v["a"]), TrackView(v["b"])
        assert (a.viewToString() == b.viewToString()) is v["expect_same"]
        assert a.viewToString() == pyref.TrackView(v["a"]).viewToString()
    assert TrackView({"level": 0, "urlId": 1}).type == "video"


def test_get_segment_time_needs_no_gpu(golden):
    mm = MediaMap(type("Hls", (), {"levels": None})())
    for v in golden["getSegmentTime"]:
        obj = {"sn": v["view"]["sn"], "trackView": {"level": v["view"]["level"], "urlId": v["view"]["urlId"]}}
        if "time" in v["view"]:
            obj["time"] = v["view"]["time"]
        if "throws" in v:
            with pytest.raises(JSError) as e:
                mm.getSegmentTime(SegmentView(obj))
            assert str(e.value) == v["throws"]
        else:
            assert mm.getSegmentTime(SegmentView(obj)) == v["expect"]
    assert mm.getTrackList() == []                       # hls.levels undefined -> []
    with pytest.raises(JSTypeError):
        mm.getSegmentList(TrackView({"level": 0, "urlId": 0}), 0, 1)


def test_js_edge_semantics():
    sv = SegmentView({"sn": 1, "trackView": {"level": 0, "urlId": 0}})
    assert sv.isEqual(None) is False and sv.isEqual(undefined) is False and sv.time is undefined
    assert sv.viewToString() == "L0U0S1" and sv.getId() == 1
    assert SegmentView({"trackView": {}}).viewToString() == "LundefinedUundefinedSundefined"
    with pytest.raises(JSTypeError):
        SegmentView({"sn": 1})                           # new TrackView(undefined)
    short = SegmentView.fromArrayBuffer(b"\x01\x00\x00\x00\x02\x00\x00\x00")
    assert (short.trackView.level, short.trackView.urlId) == (1, 2) and short.sn is undefined
    with pytest.raises(ValueError):
        SegmentView.fromArrayBuffer(b"\x00" * 6)
    assert SegmentView({"sn": -25, "trackView": {"level": 1, "urlId": 1}}).toArrayBuffer()[-4:] == b"\xe7\xff\xff\xff"


ints = st.integers(-(1 << 31), (1 << 31) - 1)


@settings(max_examples=200, deadline=None)
@given(ints, ints, ints, ints, ints, ints)
def test_mirror_equals_pyref_on_random_views(l1, u1, s1, l2, u2, s2):
    a = SegmentView({"sn": s1, "trackView": {"level": l1, "urlId": u1}, "time": 1.5})
    b = SegmentView({"sn": s2, "trackView": {"level": l2, "urlId": u2}})
    ra = pyref.SegmentView({"sn": s1, "trackView": {"level": l1, "urlId": u1}, "time": 1.5})
    rb = pyref.SegmentView({"sn": s2, "trackView": {"level": l2, "urlId": u2}})
    assert a.isEqual(b) == ra.isEqual(rb) and a.isInTrack(b.trackView) == ra.isInTrack(rb.trackView)
    assert a.viewToString() == ra.viewToString() and a.toArrayBuffer() == ra.toArrayBuffer()
    back, rback = SegmentView.fromArrayBuffer(a.toArrayBuffer()), pyref.SegmentView.fromArrayBuffer(ra.toArrayBuffer())
    assert (back.sn, back.trackView.level, back.trackView.urlId) == (rback.sn, rback.trackView.level, rback.trackView.urlId)
    assert back.isEqual(a) == rback.isEqual(ra)

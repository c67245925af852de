"""Pins the oracle (C restatement + pure-Python restatement) to the reference's own
known-answer tests: test/media-map.js, test/segment-view.js, test/track-view.js, transcribed
in tests/golden/reference_vectors.json.  CPU only."""
import json

import numpy as np
import pytest

import oracle
from oracle import pyref
from hlsjs_p2p_wrapper_b200 import FragmentTables
from hlsjs_p2p_wrapper_b200.synth import hls_mock_tables


def _mock(args):
    n, live, defined = args
    return pyref.HlsMock(n, live, defined)


def _tables_of(mock):
    lv = mock.levels
    return FragmentTables.from_hls_levels([lv if lv is not pyref.UNDEFINED else None])


def test_fixture_shape(golden):
    fx = golden["fixture"]
    m = pyref.HlsMock(3, False, 1)
    frags = m.levels[1]["details"]["fragments"]
    assert frags[0]["sn"] == fx["sn_first"] and frags[-1]["sn"] == fx["sn_last"]
    assert all(f["start"] == f["sn"] * fx["start_per_sn"] for f in frags)
    assert all("duration" not in f for f in frags)
    assert all(len(l["url"]) == fx["urls_per_level"] for l in m.levels)
    assert "details" not in m.levels[0] and "details" not in m.levels[2]
    # the SoA fixture used by the GPU tests is the same thing
    t = hls_mock_tables(3, 1, False)
    assert t.n_tracks == 3 and list(t.track_state) == [1, 2, 1]
    assert list(t.sn) == list(range(25, 200)) and list(t.start) == [10.0 * s for s in range(25, 200)]


@pytest.mark.parametrize("i", range(4))
def test_get_segment_time(golden, i):
    v = golden["getSegmentTime"][i]
    mm = pyref.MediaMap(_mock(v["mock"]))
    obj = {"sn": v["view"]["sn"], "trackView": {"level": v["view"]["level"], "urlId": v["view"]["urlId"]}}
    if "time" in v["view"]:
        obj["time"] = v["view"]["time"]
    sv = pyref.SegmentView(obj)
    if "throws" in v:
        with pytest.raises(pyref.JSError) as e:
            mm.getSegmentTime(sv)
        assert str(e.value) == v["throws"]
    else:
        assert mm.getSegmentTime(sv) == v["expect"]


def _expected(v):
    if "expect_sn_range" in v:
        sn = list(range(*v["expect_sn_range"]))
        return sn, [s * v["expect_time_per_sn"] for s in sn]
    return v["expect_sn"], v["expect_time"]


@pytest.mark.parametrize("i", range(7))
def test_get_segment_list_pyref(golden, i):
    v = golden["getSegmentList"][i]
    mm = pyref.MediaMap(_mock(v["mock"]))
    tv = pyref.TrackView(v["track"])
    if "throws" in v:
        with pytest.raises(pyref.JSError) as e:
            mm.getSegmentList(tv, v["begin"], v["duration"])
        assert str(e.value) == v["throws"]
        return
    got = mm.getSegmentList(tv, v["begin"], v["duration"])
    sn, tm = _expected(v)
    expect = [pyref.SegmentView({"sn": s, "trackView": tv, "time": t}) for s, t in zip(sn, tm)]
    assert got == expect
    assert mm.warnings == ([v["warns"]] if "warns" in v else [])


@pytest.mark.parametrize("i", range(7))
def test_get_segment_list_c_oracle(golden, i):
    v = golden["getSegmentList"][i]
    t = _tables_of(_mock(v["mock"]))
    off, status, sn, tm = oracle.get_segment_list_batch(
        t.playlist_base, t.track_off, t.track_state, t.start, t.sn,
        [0], [v["track"]["level"]], [v["begin"]], [v["duration"]])
    if "throws" in v:
        assert status[0] == oracle.ST_MISSING and off[1] == 0
        return
    esn, etm = _expected(v)
    assert status[0] == (oracle.ST_UNPARSED if "warns" in v else oracle.ST_OK)
    assert list(sn) == esn and list(tm) == [float(x) for x in etm]
    assert off[1] == len(esn)
    if status[0] == oracle.ST_OK:
        first, count, _ = oracle.get_segment_range_batch(
            t.playlist_base, t.track_off, t.track_state, t.start,
            [0], [v["track"]["level"]], [v["begin"]], [v["duration"]])
        assert count[0] == len(esn)
        assert first[0] == (esn[0] - 25 if esn else -1)


def test_get_track_list(golden):
    for v in golden["getTrackList"]:
        mock = _mock(v["mock"])
        got = pyref.MediaMap(mock).getTrackList()
        assert len(got) == v["expect_length"]
        if "expect" in v:
            assert [[t.level, t.urlId] for t in got] == v["expect"]
            t = _tables_of(mock)
            lv, ui = oracle.get_track_list(t.url_count)
            assert [[int(a), int(b)] for a, b in zip(lv, ui)] == v["expect"]


def test_segment_view_vectors(golden):
    g = golden["segmentView"]
    for v in g["json_roundtrip"]:
        sv = pyref.SegmentView(v["ctor"])
        s = sv.json_stringify()
        assert s == v["derived_json"]
        back = pyref.SegmentView(json.loads(s))
        assert back.isEqual(sv) is v["expect_equal"]
    for v in g["arraybuffer"]:
        sv = pyref.SegmentView(v["ctor"])
        buf = sv.toArrayBuffer()
        if "expect_byte_length" in v:
            assert isinstance(buf, bytes) and len(buf) == v["expect_byte_length"]
            assert buf.hex() == v["expect_bytes_hex"]
            wire = oracle.view_pack([v["ctor"]["trackView"]["level"]], [v["ctor"]["trackView"]["urlId"]],
                                    [v["ctor"]["sn"]])
            assert wire.tobytes() == buf
        if v.get("roundtrip_equal"):
            assert pyref.SegmentView.fromArrayBuffer(buf).isEqual(sv)
            lv, ui, sn = oracle.view_unpack(np.frombuffer(buf, np.uint32))
            assert (lv[0], ui[0], sn[0]) == (1, 1, 25)
            _, eq = oracle.view_roundtrip([1], [1], [25])
            assert eq[0] == 1
    for v in g["isInTrack"]:
        sv, tv = pyref.SegmentView(v["view"]), pyref.TrackView(v["track"])
        assert sv.isInTrack(tv) is v["expect"]
        assert bool(oracle.view_is_in_track([sv.trackView.level], [sv.trackView.urlId],
                                            [tv.level], [tv.urlId])[0]) is v["expect"]
    for v in g["isEqual"]:
        a, b = pyref.SegmentView(v["a"]), pyref.SegmentView(v["b"])
        assert a.isEqual(b) is v["expect"]
        assert bool(oracle.view_is_equal([a.trackView.level], [a.trackView.urlId], [a.sn],
                                         [b.trackView.level], [b.trackView.urlId], [b.sn])[0]) is v["expect"]


def test_track_view_vectors(golden):
    g = golden["trackView"]
    for v in g["isEqual"]:
        a, b = pyref.TrackView(v["a"]), pyref.TrackView(v["b"])
        assert a.isEqual(b) is v["expect"]
    for v in g["viewToString"]:
        a, b = pyref.TrackView(v["a"]), pyref.TrackView(v["b"])
        assert (a.viewToString() == b.viewToString()) is v["expect_same"]
        off, chars = oracle.view_format([a.level, b.level], [a.urlId, b.urlId], [0, 0], with_sn=False)
        sa, sb = bytes(chars[off[0]:off[1]]).decode(), bytes(chars[off[1]:off[2]]).decode()
        assert sa == a.viewToString() and sb == b.viewToString()
        assert (sa == sb) is v["expect_same"]
    assert pyref.TrackView({"level": 0, "urlId": 1}).type == "video"


def test_falsy_and_short_buffers():
    """Restated-only behaviour (no reference test): falsy argument, short / odd buffers."""
    sv = pyref.SegmentView({"sn": 1, "trackView": {"level": 0, "urlId": 0}})
    assert sv.isEqual(None) is False and sv.isEqual(pyref.UNDEFINED) is False
    assert sv.trackView.isEqual(None) is False
    short = pyref.SegmentView.fromArrayBuffer(b"\x01\x00\x00\x00\x02\x00\x00\x00")
    assert short.trackView.level == 1 and short.trackView.urlId == 2 and short.sn is pyref.UNDEFINED
    with pytest.raises(ValueError):
        pyref.SegmentView.fromArrayBuffer(b"\x00" * 7)
    with pytest.raises(pyref.JSTypeError):
        pyref.SegmentView({"sn": 1})           # new TrackView(undefined)
    assert sv.viewToString() == "L0U0S1" and sv.getId() == 1

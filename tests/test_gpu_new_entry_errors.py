"""Argument checking of the entry points added for SURVEY 8 rows f2-f4: bad calls fail with
HLSMAP_ERR_INVALID and a message, never with a crash or a silent result."""
import ctypes

import numpy as np
import pytest

from hlsjs_p2p_wrapper_b200 import synth, _native as N

pytestmark = pytest.mark.gpu


def _p(a):
    return ctypes.c_void_p(a.ctypes.data)


def test_parse_m3u8_rejects_bad_offsets(ctx):
    lib = N.load()
    text = np.frombuffer(b"#EXTINF:1,\n#EXTINF:2,\n", np.uint8)
    toff, live = np.zeros(3, np.int64), np.zeros(2, np.uint8)
    res = N.IngestResult()
    for bad in ([1, 11, 22], [0, 15, 11]):                      # does not start at 0; not monotone
        off = np.asarray(bad, np.int64)
        rc = lib.hlsmap_parse_m3u8(ctx._h, 2, None, _p(text), _p(off), _p(toff), _p(live), None, None, None, 0,
                                   ctypes.byref(res))
        assert rc == N.ERR_INVALID and lib.hlsmap_last_error()
    off = np.asarray([0, 11, 22], np.int64)
    assert lib.hlsmap_parse_m3u8(ctx._h, 2, None, None, _p(off), _p(toff), _p(live), None, None, None, 0,
                                 ctypes.byref(res)) == N.ERR_INVALID            # text missing
    assert lib.hlsmap_parse_m3u8(ctx._h, 2, None, _p(text), _p(off), None, _p(live), None, None, None, 0,
                                 ctypes.byref(res)) == N.ERR_INVALID            # no track_offset output
    assert lib.hlsmap_parse_m3u8(None, 2, None, _p(text), _p(off), _p(toff), _p(live), None, None, None, 0,
                                 ctypes.byref(res)) == N.ERR_INVALID            # no context
    # and the good call still works afterwards
    assert lib.hlsmap_parse_m3u8(ctx._h, 2, None, _p(text), _p(off), _p(toff), _p(live), None, None, None, 0,
                                 ctypes.byref(res)) == N.OK and res.n_fragments == 2 and list(toff) == [0, 1, 2]


def test_map_from_m3u8_rejects_inconsistent_tables(ctx):
    lib = N.load()
    text = np.frombuffer(b"#EXTINF:1,\n", np.uint8)
    off = np.asarray([0, 11], np.int64)
    state = np.asarray([2], np.uint8)
    h = ctypes.c_void_p()
    res = N.IngestResult()
    base_bad = np.asarray([0, 2], np.int32)                     # claims two tracks, one given
    rc = lib.hlsmap_map_create_from_m3u8(ctx._h, 1, _p(base_bad), 1, _p(state), None, _p(text), _p(off),
                                         ctypes.byref(res), ctypes.byref(h))
    assert rc == N.ERR_INVALID and not h.value
    rc = lib.hlsmap_map_create_from_m3u8(ctx._h, 1, _p(np.asarray([0, 1], np.int32)), 1, None, None, _p(text), _p(off),
                                         ctypes.byref(res), ctypes.byref(h))
    assert rc == N.ERR_INVALID and not h.value                  # track_state missing


def test_json_entry_points_reject_bad_arguments(ctx):
    lib = N.load()
    lv = np.zeros(4, np.int32)
    offs, chars, total = np.zeros(5, np.int32), np.zeros(600, np.uint8), ctypes.c_int64()
    assert lib.hlsmap_view_to_json(ctx._h, 4, _p(lv), None, _p(lv), None, _p(offs), _p(chars), 600,
                                   ctypes.byref(total)) == N.ERR_INVALID          # url_id missing
    assert lib.hlsmap_view_to_json(ctx._h, -1, _p(lv), _p(lv), _p(lv), None, _p(offs), _p(chars), 600,
                                   ctypes.byref(total)) == N.ERR_INVALID
    text = np.frombuffer(b'{"sn":1}{"sn":2}', np.uint8)
    out = [np.zeros(2, np.int32) for _ in range(3)]
    tm, fl, st = np.zeros(2), np.zeros(2, np.uint8), np.zeros(2, np.uint8)
    bad = np.asarray([0, 9, 8], np.int32)
    assert lib.hlsmap_view_from_json(ctx._h, 2, _p(text), _p(bad), _p(out[0]), _p(out[1]), _p(out[2]), _p(tm), _p(fl),
                                     _p(st)) == N.ERR_INVALID                     # offsets go backwards
    good = np.asarray([0, 8, 16], np.int32)
    assert lib.hlsmap_view_from_json(ctx._h, 2, _p(text), _p(good), _p(out[0]), _p(out[1]), _p(out[2]), _p(tm), _p(fl),
                                     None) == N.ERR_INVALID                       # status output missing
    assert lib.hlsmap_view_from_json(ctx._h, 2, _p(text), _p(good), _p(out[0]), _p(out[1]), _p(out[2]), _p(tm), _p(fl),
                                     _p(st)) == N.OK and list(out[2]) == [1, 2] and list(fl) == [1, 1]


def test_fragment_view_entry_points_reject_bad_arguments(ctx):
    lib = N.load()
    m = ctx.upload(synth.hls_mock_tables(3, 1, False))
    lv = np.zeros(3, np.int32)
    sn, tm, st = np.zeros(3, np.int32), np.zeros(3), np.zeros(3, np.uint8)
    assert lib.hlsmap_views_from_fragments(m._h, 3, None, _p(lv), None, _p(sn), _p(tm), _p(st)) == N.ERR_INVALID
    assert lib.hlsmap_views_from_fragments(None, 3, None, _p(lv), _p(lv), _p(sn), _p(tm), _p(st)) == N.ERR_INVALID
    assert lib.hlsmap_get_next_segment_view(m._h, 3, None, _p(lv), _p(lv), _p(sn), _p(tm), None, _p(st)) == N.ERR_INVALID
    m.close()

"""CPU oracle for the MediaMap path -- TEST INFRASTRUCTURE, not product code.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py`` (cpu_baseline leg and
``--impl reference``) may import this package.  It wraps ``mediamap_oracle.c`` (a plain-C
restatement of ``lib/integration/mapping/*.js``, built by ``oracle/Makefile``) with
ctypes + numpy.  ``oracle.pyref`` is the object-level pure-Python restatement.

There is no ``oracle/_ref``: the reference is JavaScript (no C/C++ sources to compile)
and no JS engine exists in the build image.  Parity is pinned on the reference's own
known-answer tests, transcribed in ``tests/golden/reference_vectors.json``.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libmediamap_oracle.so")
_lib = None

ST_OK, ST_UNPARSED, ST_MISSING, ST_EMPTY = 0, 1, 2, 3


def build(force=False):
    """Compile the C oracle with gcc (a few hundred ms)."""
    src = os.path.join(_HERE, "mediamap_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "libmediamap_oracle.so"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.oracle_filter_one.restype = ctypes.c_int64
        _lib.oracle_get_segment_list_batch.restype = ctypes.c_int64
        _lib.oracle_bench_get_segment_list.restype = ctypes.c_int64
        _lib.oracle_get_track_list.restype = ctypes.c_int64
        _lib.oracle_view_format.restype = ctypes.c_int64
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _c(a, dt):
    return None if a is None else np.ascontiguousarray(a, dtype=dt)


def _tables(playlist_base, track_off, track_state, start, sn):
    return (_c(playlist_base, np.int32), _c(track_off, np.int64), _c(track_state, np.uint8),
            _c(start, np.float64), _c(sn, np.int32))


def filter_one(start, sn, begin, duration):
    """media-map.js:42-51 on one fragment list -> (sn[], time[])."""
    start, sn = _c(start, np.float64), _c(sn, np.int32)
    so, to = np.empty(len(start), np.int32), np.empty(len(start), np.float64)
    h = lib().oracle_filter_one(_p(start), _p(sn), ctypes.c_int64(len(start)),
                                ctypes.c_double(begin), ctypes.c_double(duration), _p(so), _p(to))
    return so[:h].copy(), to[:h].copy()


def get_segment_list_batch(playlist_base, track_off, track_state, start, sn,
                           playlist, level, begin, duration, threads=1):
    """Batched list form -> (offsets int64[Q+1], status u8[Q], sn int32[H], time f64[H])."""
    pb, to, ts, st, s = _tables(playlist_base, track_off, track_state, start, sn)
    playlist, level = _c(playlist, np.int32), _c(level, np.int32)
    begin, duration = _c(begin, np.float64), _c(duration, np.float64)
    q = len(level)
    offsets = np.zeros(q + 1, np.int64)
    status = np.zeros(q, np.uint8)
    args = [ctypes.c_int32(len(pb) - 1), _p(pb), _p(to), _p(ts), _p(st), _p(s),
            ctypes.c_int64(q), _p(playlist), _p(level), _p(begin), _p(duration)]
    total = lib().oracle_get_segment_list_batch(*args, _p(offsets), _p(status), None, None,
                                                ctypes.c_int64(0), ctypes.c_int(threads))
    sn_out, time_out = np.empty(total, np.int32), np.empty(total, np.float64)
    lib().oracle_get_segment_list_batch(*args, _p(offsets), _p(status), _p(sn_out), _p(time_out),
                                        ctypes.c_int64(total), ctypes.c_int(threads))
    return offsets, status, sn_out, time_out


def get_segment_range_batch(playlist_base, track_off, track_state, start,
                            playlist, level, begin, duration):
    pb, to, ts, st, _ = _tables(playlist_base, track_off, track_state, start, None)
    playlist, level = _c(playlist, np.int32), _c(level, np.int32)
    begin, duration = _c(begin, np.float64), _c(duration, np.float64)
    q = len(level)
    first, count, status = np.empty(q, np.int32), np.empty(q, np.int32), np.empty(q, np.uint8)
    lib().oracle_get_segment_range_batch(ctypes.c_int32(len(pb) - 1), _p(pb), _p(to), _p(ts), _p(st),
                                         ctypes.c_int64(q), _p(playlist), _p(level), _p(begin),
                                         _p(duration), _p(first), _p(count), _p(status))
    return first, count, status


def bench_get_segment_list(playlist_base, track_off, track_state, start, sn,
                           playlist, level, begin, duration, threads):
    """The timed CPU baseline (one filter pass + list push per query). -> (hits, checksum)."""
    pb, to, ts, st, s = _tables(playlist_base, track_off, track_state, start, sn)
    playlist, level = _c(playlist, np.int32), _c(level, np.int32)
    begin, duration = _c(begin, np.float64), _c(duration, np.float64)
    cs = ctypes.c_uint64(0)
    hits = lib().oracle_bench_get_segment_list(
        ctypes.c_int32(len(pb) - 1), _p(pb), _p(to), _p(ts), _p(st), _p(s),
        ctypes.c_int64(len(level)), _p(playlist), _p(level), _p(begin), _p(duration),
        ctypes.c_int(threads), ctypes.byref(cs))
    return hits, cs.value


def get_segment_duration_batch(playlist_base, track_off, track_state, dur0, playlist, level):
    pb, to, ts = _c(playlist_base, np.int32), _c(track_off, np.int64), _c(track_state, np.uint8)
    dur0 = _c(dur0, np.float64)
    playlist, level = _c(playlist, np.int32), _c(level, np.int32)
    n = len(level)
    out, status = np.empty(n, np.float64), np.empty(n, np.uint8)
    lib().oracle_get_segment_duration_batch(ctypes.c_int32(len(pb) - 1), _p(pb), _p(to), _p(ts),
                                            _p(dur0), ctypes.c_int64(n), _p(playlist), _p(level),
                                            _p(out), _p(status))
    return out, status


def get_track_list(url_count):
    uc = _c(url_count, np.int32)
    n = lib().oracle_get_track_list(ctypes.c_int32(len(uc)), _p(uc), None, None)
    lv, ui = np.empty(n, np.int32), np.empty(n, np.int32)
    lib().oracle_get_track_list(ctypes.c_int32(len(uc)), _p(uc), _p(lv), _p(ui))
    return lv, ui


def view_pack(level, url_id, sn):
    level, url_id, sn = _c(level, np.int32), _c(url_id, np.int32), _c(sn, np.int32)
    wire = np.empty(3 * len(sn), np.uint32)
    lib().oracle_view_pack(ctypes.c_int64(len(sn)), _p(level), _p(url_id), _p(sn), _p(wire))
    return wire


def view_unpack(wire):
    wire = _c(wire, np.uint32)
    n = len(wire) // 3
    lv, ui, sn = np.empty(n, np.uint32), np.empty(n, np.uint32), np.empty(n, np.uint32)
    lib().oracle_view_unpack(ctypes.c_int64(n), _p(wire), _p(lv), _p(ui), _p(sn))
    return lv, ui, sn


def view_is_equal(al, au, as_, bl, bu, bs):
    arrs = [_c(x, np.int32) for x in (al, au, as_, bl, bu, bs)]
    out = np.empty(len(arrs[0]), np.uint8)
    lib().oracle_view_is_equal(ctypes.c_int64(len(out)), *[_p(a) for a in arrs], _p(out))
    return out


def view_is_in_track(al, au, tl, tu):
    arrs = [_c(x, np.int32) for x in (al, au, tl, tu)]
    out = np.empty(len(arrs[0]), np.uint8)
    lib().oracle_view_is_in_track(ctypes.c_int64(len(out)), *[_p(a) for a in arrs], _p(out))
    return out


def view_roundtrip(level, url_id, sn):
    level, url_id, sn = _c(level, np.int32), _c(url_id, np.int32), _c(sn, np.int32)
    wire, eq = np.empty(3 * len(sn), np.uint32), np.empty(len(sn), np.uint8)
    lib().oracle_view_roundtrip(ctypes.c_int64(len(sn)), _p(level), _p(url_id), _p(sn), _p(wire), _p(eq))
    return wire, eq


def view_format(level, url_id, sn, with_sn=True):
    level, url_id, sn = _c(level, np.int32), _c(url_id, np.int32), _c(sn, np.int32)
    n = len(level)
    offsets = np.empty(n + 1, np.int64)
    total = lib().oracle_view_format(ctypes.c_int64(n), _p(level), _p(url_id), _p(sn),
                                     ctypes.c_int(int(with_sn)), _p(offsets), None)
    out = np.empty(total, np.uint8)
    lib().oracle_view_format(ctypes.c_int64(n), _p(level), _p(url_id), _p(sn),
                             ctypes.c_int(int(with_sn)), _p(offsets), _p(out))
    return offsets, out


def is_monotone(start):
    start = _c(start, np.float64)
    return bool(lib().oracle_is_monotone(_p(start), ctypes.c_int64(len(start))))


def views_from_fragments(playlist_base, track_off, track_state, start, sn, playlist, level, index):
    pb, to, ts, st, s = _tables(playlist_base, track_off, track_state, start, sn)
    playlist, level, index = _c(playlist, np.int32), _c(level, np.int32), _c(index, np.int32)
    n = len(level)
    osn, tm, status = np.empty(n, np.int32), np.empty(n, np.float64), np.empty(n, np.uint8)
    lib().oracle_views_from_fragments(ctypes.c_int32(len(pb) - 1), _p(pb), _p(to), _p(ts), _p(st), _p(s),
                                      ctypes.c_int64(n), _p(playlist), _p(level), _p(index), _p(osn),
                                      _p(tm), _p(status))
    return osn, tm, status


def get_next_segment_view(playlist_base, track_off, track_state, start, sn, playlist, level, want):
    """EXTENSION restated (no reference counterpart): -> (sn, time, found, status)."""
    pb, to, ts, st, s = _tables(playlist_base, track_off, track_state, start, sn)
    playlist, level, want = _c(playlist, np.int32), _c(level, np.int32), _c(want, np.int32)
    n = len(level)
    osn, tm = np.empty(n, np.int32), np.empty(n, np.float64)
    found, status = np.empty(n, np.uint8), np.empty(n, np.uint8)
    lib().oracle_get_next_segment_view(ctypes.c_int32(len(pb) - 1), _p(pb), _p(to), _p(ts), _p(st), _p(s),
                                       ctypes.c_int64(n), _p(playlist), _p(level), _p(want), _p(osn),
                                       _p(tm), _p(found), _p(status))
    return osn, tm, found, status


def parse_m3u8(text, text_off, track_state=None, threads=1):
    """hls.js 0.5.46 parseLevelPlaylist, the part the mapping layer observes (restated; unpinned).
    -> dict(track_off, live, start, sn, duration, bad)"""
    text = np.frombuffer(bytes(text), np.uint8) if not isinstance(text, np.ndarray) else _c(text, np.uint8)
    text_off = _c(text_off, np.int64)
    state = _c(track_state, np.uint8)
    t = len(text_off) - 1
    track_off, live = np.zeros(t + 1, np.int64), np.ones(max(t, 1), np.uint8)
    fn = lib().oracle_parse_m3u8
    fn.restype = ctypes.c_int64
    n = fn(ctypes.c_int32(t), _p(state), _p(text), _p(text_off), _p(track_off), _p(live), None, None, None,
           ctypes.c_int(threads))
    bad = n < 0
    n = -1 - n if bad else n
    sn, start, dur = np.empty(n, np.int32), np.empty(n, np.float64), np.empty(n, np.float64)
    fn(ctypes.c_int32(t), _p(state), _p(text), _p(text_off), _p(track_off), _p(live), _p(sn), _p(start), _p(dur),
       ctypes.c_int(threads))
    return dict(track_off=track_off, live=live[:t], start=start, sn=sn, duration=dur, bad=bad)


def js_number_to_string(x):
    """ECMAScript Number::toString (restated with glibc printf/strtod, see mediamap_oracle.c)."""
    buf = ctypes.create_string_buffer(64)
    n = lib().oracle_js_number_to_string(ctypes.c_double(float(x)), buf)
    return buf.raw[:n].decode()


def view_to_json(level, url_id, sn, time=None):
    level, url_id, sn = _c(level, np.int32), _c(url_id, np.int32), _c(sn, np.int32)
    time = _c(time, np.float64)
    n = len(sn)
    offsets = np.zeros(n + 1, np.int32)
    fn = lib().oracle_view_to_json
    fn.restype = ctypes.c_int64
    total = fn(ctypes.c_int64(n), _p(level), _p(url_id), _p(sn), _p(time), _p(offsets), None)
    chars = np.empty(total, np.uint8)
    fn(ctypes.c_int64(n), _p(level), _p(url_id), _p(sn), _p(time), _p(offsets), _p(chars))
    return offsets, chars


def view_from_json(chars, offsets):
    chars = np.frombuffer(bytes(chars), np.uint8) if isinstance(chars, (bytes, bytearray)) else _c(chars, np.uint8)
    offsets = _c(offsets, np.int32)
    n = len(offsets) - 1
    out = dict(level=np.empty(n, np.int32), url_id=np.empty(n, np.int32), sn=np.empty(n, np.int32),
               time=np.empty(n, np.float64), flags=np.empty(n, np.uint8), status=np.empty(n, np.uint8))
    lib().oracle_view_from_json(ctypes.c_int64(n), _p(chars), _p(offsets), _p(out["level"]), _p(out["url_id"]),
                                _p(out["sn"]), _p(out["time"]), _p(out["flags"]), _p(out["status"]))
    return out

"""Host-side mirror of the reference's mapping classes, backed by the CUDA engine.

The reference's host language is JavaScript and no JS toolchain exists in this image
(see INTEGRATION.md for the N-API addon + JS drop-in that ship as source).  This module is
the same interface in Python -- same class names, method names, argument meaning, return
shapes and error messages -- so the parity tests read like the reference's mocha suite:

* :class:`MediaMap`    <- lib/integration/mapping/media-map.js:4-88
* :class:`SegmentView` <- lib/integration/mapping/segment-view.js:3-66
* :class:`TrackView`   <- lib/integration/mapping/track-view.js:1-29

What runs where (the same split the JS drop-in makes):
* ``getSegmentList`` (single call and ``getSegmentListBatch``), ``getSegmentDuration`` and
  every ``*Batch`` view operation run on the GPU through libhlsmap.so.  No CPU fallback.
* The value-type methods on ONE object (constructor, isEqual, viewToString, toArrayBuffer,
  fromArrayBuffer of one buffer, getSegmentTime, getTrackList enumeration) stay in the host
  language exactly as in the reference, because their semantics are JS object semantics
  (`undefined`, strict equality, ToUint32), not arithmetic.
"""
import struct
import warnings

import numpy as np

from . import _native as N
from .engine import Context, DeviceMap
from .tables import FragmentTables


class _Undefined:
    """JS `undefined`."""
    _inst = None

    def __new__(cls):
        if cls._inst is None:
            cls._inst = super().__new__(cls)
        return cls._inst

    def __bool__(self):
        return False

    def __repr__(self):
        return "undefined"


undefined = _Undefined()


class JSError(Exception):
    """`throw new Error(message)`; str(e) is the reference's message."""


class JSTypeError(TypeError):
    """A native TypeError (reading a property of undefined)."""


def _prop(obj, name):
    if obj is undefined or obj is None:
        raise JSTypeError("Cannot read property '%s' of %r" % (name, obj))
    if isinstance(obj, dict):
        return obj.get(name, undefined)
    return getattr(obj, name, undefined)


def _is_number(v):
    return isinstance(v, (int, float, np.integer, np.floating)) and not isinstance(v, bool)


def _same(a, b):
    """`a === b` for the values views hold."""
    if _is_number(a) and _is_number(b):
        return float(a) == float(b)
    if a is undefined or b is undefined or a is None or b is None:
        return a is b
    return type(a) is type(b) and a == b


def _js_str(v):
    if v is undefined:
        return "undefined"
    if v is None:
        return "null"
    if _is_number(v) and float(v) == int(v) and abs(float(v)) < 1e21:
        return str(int(v))
    return str(v)


def _uint32(v):
    if not _is_number(v):
        return 0
    f = float(v)
    if f != f or f in (float("inf"), float("-inf")):
        return 0
    return int(f) % (1 << 32)


class TrackView:
    def __init__(self, obj):                       # track-view.js:3-6
        self.level = _prop(obj, "level")
        self.urlId = _prop(obj, "urlId")

    def viewToString(self):                        # track-view.js:11-13
        return "L" + _js_str(self.level) + "U" + _js_str(self.urlId)

    def isEqual(self, trackView):                  # track-view.js:19-24
        if not trackView:
            return False
        return _same(_prop(trackView, "level"), self.level) and \
            _same(_prop(trackView, "urlId"), self.urlId)

    @property
    def type(self):                                # track-view.js:26-28
        return "video"

    def __eq__(self, other):
        return isinstance(other, TrackView) and (self.level, self.urlId) == (other.level, other.urlId)

    def __repr__(self):
        return "TrackView(level=%r, urlId=%r)" % (self.level, self.urlId)


class SegmentView:
    @staticmethod
    def fromArrayBuffer(arrayBuffer):              # segment-view.js:9-17
        buf = bytes(arrayBuffer)
        if len(buf) % 4:
            raise ValueError("RangeError: byte length of Uint32Array should be a multiple of 4")
        words = list(struct.unpack("<%dI" % (len(buf) // 4), buf))[:3]
        words += [undefined] * (3 - len(words))
        level, urlId, sn = words
        return SegmentView({"trackView": TrackView({"level": level, "urlId": urlId}), "sn": sn})

    def __init__(self, obj):                       # segment-view.js:22-26
        self.sn = _prop(obj, "sn")
        self.trackView = TrackView(_prop(obj, "trackView"))
        self.time = _prop(obj, "time")

    def isEqual(self, segmentView):                # segment-view.js:33-39
        if not segmentView:
            return False
        return _same(self.sn, _prop(segmentView, "sn")) and \
            self.trackView.isEqual(_prop(segmentView, "trackView"))

    def isInTrack(self, trackView):                # segment-view.js:45-47
        return self.trackView.isEqual(trackView)

    def viewToString(self):                        # segment-view.js:52-54
        return self.trackView.viewToString() + "S" + _js_str(self.sn)

    def toArrayBuffer(self):                       # segment-view.js:59-61
        return struct.pack("<3I", _uint32(self.trackView.level), _uint32(self.trackView.urlId),
                           _uint32(self.sn))

    def getId(self):                               # segment-view.js:63-65
        return self.sn

    def toJSON(self):
        """What JSON.stringify sees: own enumerable props in assignment order."""
        out = {}
        if self.sn is not undefined:
            out["sn"] = self.sn
        tv = {}
        if self.trackView.level is not undefined:
            tv["level"] = self.trackView.level
        if self.trackView.urlId is not undefined:
            tv["urlId"] = self.trackView.urlId
        out["trackView"] = tv
        if self.time is not undefined:
            out["time"] = self.time
        return out

    def __eq__(self, other):                       # should.eql-style deep equality
        return isinstance(other, SegmentView) and self.trackView == other.trackView and \
            (self.sn, self.time) == (other.sn, other.time)

    def __repr__(self):
        return "SegmentView(sn=%r, trackView=%r, time=%r)" % (self.sn, self.trackView, self.time)


_default_ctx = {}


def default_context(device=0):
    if device not in _default_ctx:
        _default_ctx[device] = Context(device)
    return _default_ctx[device]


class MediaMap:
    """`new MediaMap(hls)` (media-map.js:6-8).  `hls.levels` is re-read on every call like the
    reference does (:28,41): every call compares the level list -- every fragment's sn and start,
    every level's url count and first duration -- with what is resident in HBM.  A window that slid
    (fragments evicted at the front, appended at the back, the rest untouched) is slid in place on
    the GPU (`slide_tracks`); any other change re-uploads the tables."""

    def __init__(self, hls, context=None):
        self.hls = hls
        self._ctx = context
        self._map = None
        self._sig = None
        self._host = None          # FragmentTables resident on the GPU
        self._cap = None           # fragments reserved per track at upload (a slide cannot outgrow it)
        self.refreshes = {"upload": 0, "slide": 0}

    # -- batch construction over many players (no per-call snapshotting) --------------------
    @classmethod
    def fromTables(cls, tables: FragmentTables, context=None):
        self = cls(None, context)
        self._map = (context or default_context()).upload(tables)
        self._sig = "tables"
        return self

    def _upload(self, new):
        if self._map is not None:
            self._map.close()
        self._ctx = self._ctx or default_context()
        self._map = self._ctx.upload(new)
        self._host = new
        self._cap = np.diff(new.track_off).astype(np.int64)
        self.refreshes["upload"] += 1

    def _device_map(self):
        if self._sig == "tables":
            return self._map
        new = FragmentTables.from_hls_levels([self.hls.levels])
        old = self._host
        if old is None or new.n_tracks != old.n_tracks or \
                not np.array_equal(new.track_state, old.track_state) or \
                not np.array_equal(new.url_count, old.url_count):
            self._upload(new)
            return self._map
        bits = lambda a: a.view(np.uint64)
        slides = []
        for t in range(new.n_tracks):
            o0, o1, n0, n1 = int(old.track_off[t]), int(old.track_off[t + 1]), int(new.track_off[t]), int(new.track_off[t + 1])
            osn, ost, nsn, nst = old.sn[o0:o1], old.start[o0:o1], new.sn[n0:n1], new.start[n0:n1]
            d_old = bits(old.duration[o0:o0 + 1]) if o1 > o0 else None
            d_new = bits(new.duration[n0:n0 + 1]) if n1 > n0 else None
            same_dur0 = (d_old is None and d_new is None) or \
                (d_old is not None and d_new is not None and d_old[0] == d_new[0])
            if len(osn) == len(nsn) and np.array_equal(osn, nsn) and np.array_equal(bits(ost), bits(nst)) and same_dur0:
                continue
            # a slid window?  old[k:] must be the head of new
            k = None
            if len(nsn) and len(osn):
                hit = np.nonzero((osn == nsn[0]) & (bits(ost) == bits(nst)[0]))[0]
                k = int(hit[0]) if len(hit) else None
            elif len(nsn) == 0:
                k = len(osn)
            keep = len(osn) - k if k is not None else -1
            ok = k is not None and keep <= len(nsn) and len(nsn) <= self._cap[t] and \
                np.array_equal(osn[k:], nsn[:keep]) and np.array_equal(bits(ost[k:]), bits(nst[:keep]))
            if not ok:
                self._upload(new)
                return self._map
            slides.append((t, k, nst[keep:], nsn[keep:], new.duration[n0] if n1 > n0 else np.nan))
        if slides:
            aoff = np.concatenate([[0], np.cumsum([len(x[2]) for x in slides])]).astype(np.int64)
            self._map.slide_tracks(np.zeros(len(slides), np.int32), np.asarray([x[0] for x in slides], np.int32),
                                   np.asarray([x[1] for x in slides], np.int32), aoff,
                                   np.concatenate([x[2] for x in slides]) if aoff[-1] else np.zeros(0),
                                   np.concatenate([x[3] for x in slides]).astype(np.int32) if aoff[-1] else np.zeros(0, np.int32),
                                   new_dur0=np.asarray([x[4] for x in slides], np.float64))
            self._host = new
            self.refreshes["slide"] += 1
        return self._map

    def getSegmentTime(self, segmentView):         # media-map.js:14-19
        if _prop(segmentView, "time") is undefined:
            raise JSError("getSegmentTime: segmentView.time is undefined")
        return segmentView.time

    def getSegmentList(self, trackView, beginTime, duration):   # media-map.js:27-54
        levels = self.hls.levels
        if levels is undefined or levels is None:
            raise JSTypeError("Cannot read property '%s' of undefined" % _js_str(_prop(trackView, "level")))
        lvl = _prop(trackView, "level")
        if not (_is_number(lvl) and float(lvl) == int(lvl) and 0 <= int(lvl) < (1 << 31)):
            # levels[<non-index>] is undefined in JS
            raise JSError("getSegmentList: level doesn't exist")
        res = self._device_map().get_segment_list(
            np.zeros(1, np.int32), np.asarray([int(lvl)], np.int32),
            np.asarray([beginTime], np.float64), np.asarray([duration], np.float64))
        if res.status[0] == N.ST_MISSING:
            raise JSError("getSegmentList: level doesn't exist")
        if res.status[0] == N.ST_UNPARSED:
            warnings.warn("getSegmentList: level not parsed yet")        # console.warn (:35)
            return []
        return [SegmentView({"sn": int(s), "trackView": trackView, "time": float(t)})
                for s, t in zip(res.sn, res.time)]

    def getSegmentListBatch(self, level, beginTime, duration, playlist=None, strict=True):
        """The added batched entry point: typed arrays in, typed arrays out, no per-hit
        objects.  Returns a SegmentListResult (offsets[Q+1], sn[H], time[H], status[Q]).
        With strict=True a query on a non-existent level raises the reference's error and
        an unparsed level warns once, as Q single calls would."""
        level = np.asarray(level, np.int32)
        pl = np.zeros(len(level), np.int32) if playlist is None else np.asarray(playlist, np.int32)
        res = self._device_map().get_segment_list(pl, level, beginTime, duration)
        if strict and res.n_missing:
            raise JSError("getSegmentList: level doesn't exist")
        if strict and res.n_unparsed:
            warnings.warn("getSegmentList: level not parsed yet")
        return res

    def getTrackList(self):                        # media-map.js:60-73
        if self._sig != "tables" and not self.hls.levels:
            return []
        if self._sig != "tables":
            for lv in self.hls.levels:             # `level.url.length` (:67-68) throws on a falsy level
                if not lv:
                    raise JSTypeError("Cannot read property 'url' of undefined")
                if _prop(lv, "url") is undefined or _prop(lv, "url") is None:
                    raise JSTypeError("Cannot read property 'length' of undefined")
        lv, ui = self._device_map().get_track_list(0)
        return [TrackView({"level": int(a), "urlId": int(b)}) for a, b in zip(lv, ui)]

    def getSegmentDuration(self, segmentView):     # media-map.js:81-87
        lvl = _prop(_prop(segmentView, "trackView"), "level")
        if self._sig != "tables":
            levels = self.hls.levels
            level = levels[lvl] if _is_number(lvl) and 0 <= int(lvl) < len(levels) else undefined
            details = _prop(level, "details")      # TypeError on a missing level, like JS
            _prop(details, "fragments")            # TypeError on an unparsed level
        out, status = self._device_map().get_segment_duration(np.zeros(1, np.int32),
                                                              np.asarray([int(lvl)], np.int32))
        if status[0] == N.ST_EMPTY:
            raise JSError("All segments should have a duration")
        if status[0] != N.ST_OK:
            raise JSTypeError("Cannot read property 'details' of undefined")
        v = float(out[0])
        return undefined if v != v else v          # a fragment without `duration` reads undefined

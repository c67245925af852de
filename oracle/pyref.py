"""Pure-Python restatement of the reference's mapping layer.  TEST INFRASTRUCTURE ONLY.

Only tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline leg may import
this module; the product package never does.  It restates, object for object, what the
reference's JavaScript does so that small cases (and the C oracle in
``mediamap_oracle.c``) can be checked against something that reads like the source:

* ``lib/integration/mapping/media-map.js:6-87``   -> :class:`MediaMap`
* ``lib/integration/mapping/segment-view.js:9-65`` -> :class:`SegmentView`
* ``lib/integration/mapping/track-view.js:3-28``  -> :class:`TrackView`
* ``test/mocks/hls.js:5-45``                       -> :class:`HlsMock` (fixture, without
  its ``hls.js`` package import, which only serves ``get config()``)

Python floats are IEEE-754 binary64 like JS Numbers, so ``beginTime + duration`` rounds
identically.  JS ``undefined`` is modelled by :data:`UNDEFINED`.

Parity pin: tests/test_oracle_golden.py runs every known-answer vector of the reference's
mocha suite (tests/golden/reference_vectors.json) through this module and the C oracle.
"""
import json
import struct


class _Undefined:
    """JS `undefined`: falsy, equal only to itself, dropped by JSON.stringify."""
    _inst = None

    def __new__(cls):
        if cls._inst is None:
            cls._inst = super().__new__(cls)
        return cls._inst

    def __bool__(self):
        return False

    def __repr__(self):
        return "undefined"


UNDEFINED = _Undefined()


class JSError(Exception):
    """`throw new Error(msg)`."""


class JSTypeError(Exception):
    """Native TypeError (property access on undefined)."""


def _get(obj, key):
    """`obj.key` with JS semantics: TypeError on undefined/null, undefined if absent."""
    if obj is UNDEFINED or obj is None:
        raise JSTypeError("Cannot read property '%s' of %r" % (key, obj))
    if isinstance(obj, dict):
        return obj.get(key, UNDEFINED)
    return getattr(obj, key, UNDEFINED)


def _strict_eq(a, b):
    """JS `===` on the value kinds that occur here (numbers, undefined, null, strings)."""
    if a is UNDEFINED or b is UNDEFINED or a is None or b is None:
        return a is b
    if isinstance(a, bool) != isinstance(b, bool):
        return False
    if isinstance(a, (int, float)) and isinstance(b, (int, float)):
        return float(a) == float(b)
    return type(a) is type(b) and a == b


def _num_to_string(v):
    """JS Number -> String for the values `viewToString` sees in the batch domain."""
    if v is UNDEFINED:
        return "undefined"
    if v is None:
        return "null"
    if isinstance(v, float) and v == int(v) and abs(v) < 1e21:
        return str(int(v))
    return str(v)


def _to_uint32(v):
    """ECMAScript ToUint32 (what `new Uint32Array([...])` applies, segment-view.js:60)."""
    if v is UNDEFINED or v is None:
        return 0
    f = float(v)
    if f != f or f in (float("inf"), float("-inf")):
        return 0
    return int(f) % (1 << 32) if f >= 0 else (-(int(-f))) % (1 << 32)


class TrackView:
    # track-view.js:3-6
    def __init__(self, obj):
        self.level = _get(obj, "level")
        self.urlId = _get(obj, "urlId")

    # track-view.js:11-13
    def viewToString(self):
        return "L%sU%s" % (_num_to_string(self.level), _num_to_string(self.urlId))

    # track-view.js:19-24
    def isEqual(self, trackView):
        if not trackView:
            return False
        return _strict_eq(_get(trackView, "level"), self.level) and \
            _strict_eq(_get(trackView, "urlId"), self.urlId)

    # track-view.js:26-28
    @property
    def type(self):
        return "video"

    def to_json_obj(self):
        return {k: v for k, v in (("level", self.level), ("urlId", self.urlId))
                if v is not UNDEFINED}


class SegmentView:
    # segment-view.js:9-17
    @staticmethod
    def fromArrayBuffer(arrayBuffer):
        buf = bytes(arrayBuffer)
        if len(buf) % 4:
            raise ValueError("RangeError: byte length of Uint32Array should be a multiple of 4")
        u32 = struct.unpack("<%dI" % (len(buf) // 4), buf)
        vals = list(u32[:3]) + [UNDEFINED] * (3 - min(3, len(u32)))
        level, urlId, sn = vals
        return SegmentView({"trackView": TrackView({"level": level, "urlId": urlId}), "sn": sn})

    # segment-view.js:22-26
    def __init__(self, obj):
        self.sn = _get(obj, "sn")
        self.trackView = TrackView(_get(obj, "trackView"))
        self.time = _get(obj, "time")

    # segment-view.js:33-39
    def isEqual(self, segmentView):
        if not segmentView:
            return False
        sn, trackView = _get(segmentView, "sn"), _get(segmentView, "trackView")
        return _strict_eq(self.sn, sn) and self.trackView.isEqual(trackView)

    # segment-view.js:45-47
    def isInTrack(self, trackView):
        return self.trackView.isEqual(trackView)

    # segment-view.js:52-54
    def viewToString(self):
        return "%sS%s" % (self.trackView.viewToString(), _num_to_string(self.sn))

    # segment-view.js:59-61
    def toArrayBuffer(self):
        return struct.pack("<3I", _to_uint32(self.trackView.level),
                           _to_uint32(self.trackView.urlId), _to_uint32(self.sn))

    # segment-view.js:63-65
    def getId(self):
        return self.sn

    def to_json_obj(self):
        """Own enumerable properties in assignment order (sn, trackView, time)."""
        out = {}
        if self.sn is not UNDEFINED:
            out["sn"] = self.sn
        out["trackView"] = self.trackView.to_json_obj()
        if self.time is not UNDEFINED:
            out["time"] = self.time
        return out

    def json_stringify(self):
        return json.dumps(self.to_json_obj(), separators=(",", ":"))

    def __eq__(self, other):
        """should.eql deep equality (test/media-map.js:57): same class, same own props."""
        return isinstance(other, SegmentView) and \
            (self.sn, self.trackView.level, self.trackView.urlId, self.time) == \
            (other.sn, other.trackView.level, other.trackView.urlId, other.time)

    def __repr__(self):
        return "SegmentView(sn=%r, level=%r, urlId=%r, time=%r)" % (
            self.sn, self.trackView.level, self.trackView.urlId, self.time)


class MediaMap:
    # media-map.js:6-8
    def __init__(self, hls):
        self.hls = hls
        self.warnings = []          # stands in for console.warn (media-map.js:35)

    # media-map.js:14-19
    def getSegmentTime(self, segmentView):
        if _get(segmentView, "time") is UNDEFINED:
            raise JSError("getSegmentTime: segmentView.time is undefined")
        return segmentView.time

    # media-map.js:27-54
    def getSegmentList(self, trackView, beginTime, duration):
        levels = self.hls.levels
        if levels is UNDEFINED or levels is None:
            raise JSTypeError("Cannot read property of undefined")
        idx = _get(trackView, "level")
        level = levels[idx] if isinstance(idx, int) and 0 <= idx < len(levels) else UNDEFINED
        if not level:
            raise JSError("getSegmentList: level doesn't exist")
        details = _get(level, "details")
        if not details:
            self.warnings.append("getSegmentList: level not parsed yet")
            return []
        segmentList = []
        fragments = _get(details, "fragments")
        for fragment in fragments:
            start = _get(fragment, "start")
            if beginTime <= start and start <= beginTime + duration:
                segmentList.append(SegmentView({"sn": _get(fragment, "sn"),
                                                "trackView": trackView, "time": start}))
        return segmentList

    # media-map.js:60-73
    def getTrackList(self):
        if not self.hls.levels:
            return []
        trackList = []
        for i, level in enumerate(self.hls.levels):
            for j in range(len(_get(level, "url"))):
                trackList.append(TrackView({"level": i, "urlId": j}))
        return trackList

    # media-map.js:81-87
    def getSegmentDuration(self, segmentView):
        level = self.hls.levels[segmentView.trackView.level]
        for fragment in _get(_get(level, "details"), "fragments"):
            return _get(fragment, "duration")
        raise JSError("All segments should have a duration")


class HlsMock:
    """test/mocks/hls.js:5-45 restated: 175 fragments sn 25..199, start = 10*sn, no
    `duration` field; two URLs per level; only `definedLevel` gets details."""

    def __init__(self, levelNumber, live, definedLevel=0, emptyLevel=True):
        self._levels = UNDEFINED
        if levelNumber > 0:
            self._levels = []
        fragments = [{"sn": f, "start": f * 10} for f in range(25, 200)]
        for i in range(levelNumber):
            url = ["http://foo.bar/%d/0/playlist.m3u8" % i, "http://foo.bar/%d/1/playlist.m3u8" % i]
            if emptyLevel:
                level = {"url": url}
            else:
                level = {"details": {"totalduration": 120}, "audioCodec": "fooCodec", "url": url}
            if live is not UNDEFINED and i == definedLevel:
                level["details"] = {"live": live, "fragments": fragments}
            self._levels.append(level)

    @property
    def levels(self):
        return self._levels


# ---- hls.js 0.5.46 parseLevelPlaylist, the part the mapping layer observes ----
# The dependency is not in the reference checkout (package.json:34), so this is the published
# expression as restated for this build, run by a real regular-expression engine: an independent
# formulation of what oracle_parse_m3u8 (mediamap_oracle.c) emulates by hand.  Parity unpinned.
import re as _re

_JS_DOT = "[^\\n\\r\\u2028\\u2029]"          # JavaScript's `.`; Python's would also take \r
LEVEL_PLAYLIST_RE = _re.compile(
    "(?:#EXT-X-(MEDIA-SEQUENCE):(\\d+))|(?:#EXT-X-(TARGETDURATION):(\\d+))|(?:#EXT-X-(KEY):(" + _JS_DOT + "*))"
    "|(?:#EXT(INF):([\\d\\.]+)[^\\r\\n]*([\\r\\n]+[^#|\\r\\n]+)?)"
    "|(?:#EXT-X-(BYTERANGE):([\\d]+[@\\[\\d]*)\\]*[\\r\\n]+([^#|\\r\\n]+)?|(?:#EXT-X-(ENDLIST))|(?:#EXT-X-(DIS)CONTINUITY))"
    "|(?:#EXT-X-(PROGRAM-DATE-TIME):(" + _JS_DOT + "*))", _re.ASCII)
_STR_DECIMAL = _re.compile(r"(?:\d+\.?\d*|\.\d+)", _re.ASCII)


def js_parse_float_of_run(run):
    """parseFloat on a [0-9.]+ run: the longest StrDecimalLiteral prefix, else NaN."""
    m = _STR_DECIMAL.match(run)
    return float(m.group(0)) if m else float("nan")


def parse_level_playlist(text):
    """-> (fragments [{sn, start, duration}], live).  `text` is a str (one level playlist)."""
    current_sn, total, live, frags = 0, 0.0, True, []
    for m in LEVEL_PLAYLIST_RE.finditer(text):
        result = [g for g in m.groups() if g is not None]       # result.shift(); filter(undefined)
        tag = result[0]
        if tag == "MEDIA-SEQUENCE":
            current_sn = int(result[1])
        elif tag == "ENDLIST":
            live = False
        elif tag == "INF":
            duration = js_parse_float_of_run(result[1])
            if duration == duration:
                frags.append({"sn": current_sn, "start": total, "duration": duration})
                current_sn += 1
                total += duration
    return frags, live

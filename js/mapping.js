'use strict';
/*
 * Drop-in for streamroot-hlsjs-p2p-wrapper's lib/integration/mapping/{media-map,segment-view,
 * track-view}.js backed by libhlsmap (B200).  Same constructors, methods, return shapes and
 * error messages; `getSegmentList` runs on the GPU through the N-API addon (js/addon), and
 * MediaMap gains `getSegmentListBatch`.  One file, CommonJS; `module.exports.MediaMap` etc. are
 * what the three reference modules default-export.
 *
 * Not runnable in the build image (no Node there); the same behaviour is exercised through the
 * Python mirror hlsjs-p2p-wrapper_b200/mapping.py by tests/test_gpu_media_map.py.
 */
const native = require('./addon/build/Release/hlsmap_napi.node');

// ---- value types (host side: these are JS object semantics, not arithmetic) -----------------
function TrackView(fields) {            // track-view.js:3-6
  this.level = fields.level;
  this.urlId = fields.urlId;
}
TrackView.prototype.viewToString = function () {            // track-view.js:11-13
  return 'L' + this.level + 'U' + this.urlId;
};
TrackView.prototype.isEqual = function (other) {            // track-view.js:19-24
  return !!other && other.level === this.level && other.urlId === this.urlId;
};
Object.defineProperty(TrackView.prototype, 'type', {        // track-view.js:26-28
  get() { return 'video'; },
});

function SegmentView(fields) {          // segment-view.js:22-26 (key order = JSON order)
  this.sn = fields.sn;
  this.trackView = new TrackView(fields.trackView);
  this.time = fields.time;
}
SegmentView.fromArrayBuffer = function (buffer) {           // segment-view.js:9-17
  const words = new Uint32Array(buffer);
  return new SegmentView({ trackView: new TrackView({ level: words[0], urlId: words[1] }), sn: words[2] });
};
SegmentView.prototype.isEqual = function (other) {          // segment-view.js:33-39
  return !!other && this.sn === other.sn && this.trackView.isEqual(other.trackView);
};
SegmentView.prototype.isInTrack = function (trackView) {    // segment-view.js:45-47
  return this.trackView.isEqual(trackView);
};
SegmentView.prototype.viewToString = function () {          // segment-view.js:52-54
  return this.trackView.viewToString() + 'S' + this.sn;
};
SegmentView.prototype.toArrayBuffer = function () {         // segment-view.js:59-61
  return Uint32Array.of(this.trackView.level, this.trackView.urlId, this.sn).buffer;
};
SegmentView.prototype.getId = function () { return this.sn; };   // segment-view.js:63-65

// Batched forms over SoA typed arrays: GPU kernels (views.cu)
SegmentView.packBatch = (ctx, level, urlId, sn) => native.viewPack(ctx, level, urlId, sn);
SegmentView.unpackBatch = (ctx, wire) => native.viewUnpack(ctx, wire);
SegmentView.isEqualBatch = (ctx, a, b) => native.viewIsEqual(ctx, a.level, a.urlId, a.sn, b.level, b.urlId, b.sn);
SegmentView.viewToStringBatch = (ctx, level, urlId, sn) => native.viewFormat(ctx, level, urlId, sn, true);
SegmentView.isInTrackBatch = (ctx, a, track) => native.viewIsInTrack(ctx, a.level, a.urlId, track.level, track.urlId);
// toArrayBuffer -> fromArrayBuffer -> isEqual(original) in one pass: {wire: Uint32Array(3n), equal: Uint8Array}
SegmentView.roundtripBatch = (ctx, level, urlId, sn) => native.viewRoundtrip(ctx, level, urlId, sn);

// ---- MediaMap --------------------------------------------------------------------------------
const ST_UNPARSED = 1, ST_MISSING = 2, ST_EMPTY = 3;
let sharedContext = null;

function snapshotLevels(levels) {
  // hls.levels[..].details.fragments (AoS of JS objects) -> SoA typed arrays (a13)
  const nTracks = levels.length;
  let total = 0;
  for (const level of levels) if (level && level.details) total += level.details.fragments.length;
  const t = {
    playlistBase: Int32Array.of(0, nTracks), trackOffset: new BigInt64Array(nTracks + 1),
    trackState: new Uint8Array(nTracks), urlCount: new Int32Array(nTracks),
    start: new Float64Array(total), sn: new Int32Array(total), duration: new Float64Array(total),
  };
  let at = 0;
  levels.forEach((level, i) => {
    if (level) {
      t.urlCount[i] = level.url ? level.url.length : 0;
      t.trackState[i] = level.details ? 2 : 1;
      if (level.details) {
        for (const f of level.details.fragments) {
          t.start[at] = f.start; t.sn[at] = f.sn;
          t.duration[at] = f.duration === undefined ? NaN : f.duration;
          at++;
        }
      }
    }
    t.trackOffset[i + 1] = BigInt(at);
  });
  return t;
}

// Bit pattern view of a Float64Array: NaN starts compare equal to themselves, -0 differs from 0.
const bitsOf = (f64) => new BigUint64Array(f64.buffer, f64.byteOffset, f64.length);

function sameRun(aSn, aBits, a0, bSn, bBits, b0, n) {
  for (let i = 0; i < n; i++) {
    if (aSn[a0 + i] !== bSn[b0 + i] || aBits[a0 + i] !== bBits[b0 + i]) return false;
  }
  return true;
}

class MediaMap {
  constructor(hls, options) {           // media-map.js:6-8
    this.hls = hls;
    this._device = options && options.device !== undefined ? options.device : 0;
    this._map = null;
    this._host = null;                  // the tables resident on the GPU (snapshotLevels form)
    this._cap = null;                   // fragments reserved per level at upload
    this.refreshes = { upload: 0, slide: 0 };
  }

  _upload(t) {
    if (!sharedContext) sharedContext = native.createContext(this._device);
    if (this._map) native.destroyMap(this._map);
    this._map = native.createMap(sharedContext, t);
    this._host = t;
    this._cap = Array.from({ length: t.trackState.length }, (_, i) => Number(t.trackOffset[i + 1] - t.trackOffset[i]));
    this.refreshes.upload++;
    return this._map;
  }

  // The reference reads hls.levels on every call (:28,41).  Every call therefore compares the level
  // list -- every fragment's sn and start, every level's url count and first duration -- with what is
  // resident in HBM.  A window that slid (fragments evicted at the front, appended at the back, the
  // rest untouched) is slid in place on the GPU (native.slideTracks); anything else re-uploads.
  _deviceMap() {
    const t = snapshotLevels(this.hls.levels);
    const old = this._host;
    const nT = t.trackState.length;
    if (!old || old.trackState.length !== nT) return this._upload(t);
    for (let i = 0; i < nT; i++) {
      if (t.trackState[i] !== old.trackState[i] || t.urlCount[i] !== old.urlCount[i]) return this._upload(t);
    }
    const nb = bitsOf(t.start), ob = bitsOf(old.start), nd = bitsOf(t.duration), od = bitsOf(old.duration);
    const slides = [];
    for (let i = 0; i < nT; i++) {
      const o0 = Number(old.trackOffset[i]), oN = Number(old.trackOffset[i + 1]) - o0;
      const n0 = Number(t.trackOffset[i]), nN = Number(t.trackOffset[i + 1]) - n0;
      const sameDur0 = (oN === 0 && nN === 0) || (oN > 0 && nN > 0 && od[o0] === nd[n0]);
      if (oN === nN && sameDur0 && sameRun(old.sn, ob, o0, t.sn, nb, n0, nN)) continue;
      let k = -1;                        // old[k..] must be the head of new
      if (nN === 0) k = oN;
      else for (let j = 0; j < oN; j++) if (old.sn[o0 + j] === t.sn[n0] && ob[o0 + j] === nb[n0]) { k = j; break; }
      const keep = oN - k;
      if (k < 0 || keep > nN || nN > this._cap[i] || !sameRun(old.sn, ob, o0 + k, t.sn, nb, n0, keep)) return this._upload(t);
      slides.push({ level: i, drop: k, from: n0 + keep, to: n0 + nN, dur0: nN > 0 ? t.duration[n0] : NaN });
    }
    if (slides.length) {
      const n = slides.length;
      const appendOffset = new BigInt64Array(n + 1);
      slides.forEach((s, j) => { appendOffset[j + 1] = appendOffset[j] + BigInt(s.to - s.from); });
      const total = Number(appendOffset[n]);
      const newStart = new Float64Array(total), newSn = new Int32Array(total);
      slides.forEach((s, j) => {
        newStart.set(t.start.subarray(s.from, s.to), Number(appendOffset[j]));
        newSn.set(t.sn.subarray(s.from, s.to), Number(appendOffset[j]));
      });
      native.slideTracks(this._map, {
        playlist: null, level: Int32Array.from(slides, (s) => s.level), drop: Int32Array.from(slides, (s) => s.drop),
        appendOffset, newStart, newSn, newDur0: Float64Array.from(slides, (s) => s.dur0),
      });
      this._host = t;
      this.refreshes.slide++;
    }
    return this._map;
  }

  getSegmentTime(segmentView) {         // media-map.js:14-19
    if (segmentView.time === undefined) throw new Error('getSegmentTime: segmentView.time is undefined');
    return segmentView.time;
  }

  getSegmentList(trackView, beginTime, duration) {          // media-map.js:27-54
    const levels = this.hls.levels;     // TypeError below if undefined, like the reference
    if (!levels[trackView.level]) throw new Error("getSegmentList: level doesn't exist");
    if (!levels[trackView.level].details) {
      console.warn('getSegmentList: level not parsed yet');
      return [];
    }
    const r = native.getSegmentList(this._deviceMap(), null, Int32Array.of(trackView.level),
      Float64Array.of(beginTime), Float64Array.of(duration));
    const out = new Array(r.totalHits);
    for (let i = 0; i < r.totalHits; i++) out[i] = new SegmentView({ sn: r.sn[i], trackView, time: r.time[i] });
    return out;
  }

  // Added entry point: typed arrays in, typed arrays out, no per-hit objects.
  // queries = {playlist?: Int32Array, level: Int32Array, beginTime: Float64Array, duration: Float64Array}
  // -> {offsets: Int32Array(Q+1), sn: Int32Array(H), time: Float64Array(H), status: Uint8Array(Q)}
  // Range form of the same query: {first: Int32Array(Q), count: Int32Array(Q), status: Uint8Array(Q)} --
  // fragment index of the first kept fragment (or -1) and how many follow, 9 bytes back per query.
  getSegmentRangeBatch(queries) {
    return native.getSegmentRange(this._deviceMap(), queries.playlist || null, queries.level,
      queries.beginTime, queries.duration);
  }

  getSegmentListBatch(queries, options) {
    const r = native.getSegmentList(this._deviceMap(), queries.playlist || null, queries.level,
      queries.beginTime, queries.duration);
    if (!(options && options.strict === false)) {
      if (r.nMissing > 0) throw new Error("getSegmentList: level doesn't exist");
      if (r.nUnparsed > 0) console.warn('getSegmentList: level not parsed yet');
    }
    return r;
  }

  getTrackList() {                      // media-map.js:60-73
    if (!this.hls.levels) return [];
    for (const level of this.hls.levels) void level.url.length;   // a falsy level throws a TypeError here (:67-68)
    const r = native.getTrackList(this._deviceMap(), 0);
    return Array.from(r.level, (level, i) => new TrackView({ level, urlId: r.urlId[i] }));
  }

  getSegmentDuration(segmentView) {     // media-map.js:81-87
    const level = this.hls.levels[segmentView.trackView.level];
    void level.details.fragments;       // TypeError on a missing / unparsed level, like the reference
    const r = native.getSegmentDuration(this._deviceMap(), null, Int32Array.of(segmentView.trackView.level));
    if (r.status[0] === ST_EMPTY) throw new Error('All segments should have a duration');
    return Number.isNaN(r.out[0]) ? undefined : r.out[0];
  }
}

// ---- batch forms either side of the path (no counterpart in the reference's classes) ----------
// The level playlists of many players, still as M3U8 text (what hls.js 0.5.46 would have parsed,
// package.json:34), parsed on the GPU straight into a map.  tables = {playlistBase: Int32Array(P+1),
// trackState: Uint8Array(T), urlCount: Int32Array(T)|null, text: Uint8Array, textOffset: BigInt64Array(T+1)}
function createMapFromM3u8(tables, device) {
  if (!sharedContext) sharedContext = native.createContext(device === undefined ? 0 : device);
  return native.createMapFromM3u8(sharedContext, tables);
}

// JSON.stringify of many SegmentViews at once (test/segment-view.js:5-11, batched): columns in,
// {offsets: Int32Array(n+1), chars: Uint8Array} out; texts[i] = chars.subarray(offsets[i], offsets[i+1]).
function segmentViewsToJSON(columns) {
  if (!sharedContext) sharedContext = native.createContext(0);
  return native.viewToJson(sharedContext, columns.level, columns.urlId, columns.sn, columns.time || null);
}

// ... and back: {level, urlId, sn, time, flags, status} typed arrays, one entry per text.
function segmentViewsFromJSON(chars, offsets) {
  if (!sharedContext) sharedContext = native.createContext(0);
  return native.viewFromJson(sharedContext, chars, offsets);
}

// Typed arrays over pinned host memory (hlsmap_host_alloc): batches built in them reach the GPU
// without a staging copy.  pinned(Float64Array, n) etc.; the memory is released with the ArrayBuffer.
function pinned(TypedArray, length, device) {
  if (!sharedContext) sharedContext = native.createContext(device === undefined ? 0 : device);
  return new TypedArray(native.hostAlloc(sharedContext, length * TypedArray.BYTES_PER_ELEMENT), 0, length);
}

// Consecutive getSegmentListBatch calls that share no buffers may overlap on the GPU.
function setBatchOverlap(enable) {
  if (!sharedContext) sharedContext = native.createContext(0);
  native.setBatchOverlap(sharedContext, !!enable);
}

module.exports = {
  MediaMap, SegmentView, TrackView, native, ST_UNPARSED, ST_MISSING, ST_EMPTY,
  createMapFromM3u8, segmentViewsToJSON, segmentViewsFromJSON, pinned, setBatchOverlap,
};

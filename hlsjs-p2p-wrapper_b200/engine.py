"""numpy-facing wrapper of the C ABI: one :class:`Context` per GPU, one :class:`DeviceMap`
per uploaded set of playlists.  Everything computes on the GPU through libhlsmap.so; the
``*_dev`` methods take raw device pointers (ints, e.g. ``tensor.data_ptr()``) and are what
the benchmark times."""
import ctypes

import numpy as np

from . import _native as N
from .tables import FragmentTables


def _ptr(a):
    return None if a is None else ctypes.c_void_p(a.ctypes.data)


def _arr(a, dtype):
    return None if a is None else np.ascontiguousarray(a, dtype=dtype)


CUDA_STREAM_LEGACY = 1   # the cudaStreamLegacy handle


def _stream(s):
    """None -> the context's own stream; 0 -> CUDA's legacy default stream (what
    `torch.cuda.default_stream().cuda_stream` is); anything else is a cudaStream_t."""
    if s is None:
        return None
    return ctypes.c_void_p(CUDA_STREAM_LEGACY if int(s) == 0 else int(s))


class SegmentListResult:
    """List form of a getSegmentList batch (media-map.js:27-54 per query)."""
    __slots__ = ("offsets", "sn", "time", "status", "total_hits", "n_unparsed", "n_missing",
                 "kernel_launches", "path")

    def hits(self, q):
        a, b = int(self.offsets[q]), int(self.offsets[q + 1])
        return self.sn[a:b], self.time[a:b]


class Context:
    def __init__(self, device=0):
        self._lib = N.load()
        h = ctypes.c_void_p()
        N.check(self._lib.hlsmap_ctx_create(int(device), ctypes.byref(h)))
        self._h = h
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.hlsmap_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_batch_overlap(self, enable=True):
        """Allow consecutive, independent device-pointer batches to overlap (the next one starts
        while the previous one drains).  The caller guarantees they share no output buffers.
        enable=2: overlapping launches, but a batches call keeps one launch per batch (no fusing)."""
        N.check(self._lib.hlsmap_ctx_set_batch_overlap(self._h, 2 if enable == 2 else int(bool(enable))))

    def synchronize(self):
        N.check(self._lib.hlsmap_ctx_synchronize(self._h))

    def upload(self, tables: FragmentTables):
        return DeviceMap(self, tables)

    # ---- JSON form of views ---------------------------------------------
    def view_to_json(self, level, url_id, sn, time=None):
        """JSON.stringify of n SegmentViews -> (offsets int32[n+1], chars uint8[total])."""
        level, url_id, sn = _arr(level, np.int32), _arr(url_id, np.int32), _arr(sn, np.int32)
        time = None if time is None else _arr(time, np.float64)
        n = len(sn)
        offsets = np.zeros(n + 1, np.int32)
        chars = np.empty(max(104 * n, 16), np.uint8)
        total = ctypes.c_int64(0)
        N.check(self._lib.hlsmap_view_to_json(self._h, n, _ptr(level), _ptr(url_id), _ptr(sn), _ptr(time),
                                              _ptr(offsets), _ptr(chars), len(chars), ctypes.byref(total)))
        return offsets, chars[:total.value].copy()

    def view_from_json(self, chars, offsets):
        """Texts in the shape JSON.stringify(segmentView) gives -> dict of columns, flags, status."""
        chars = np.frombuffer(chars, np.uint8) if isinstance(chars, (bytes, bytearray)) else _arr(chars, np.uint8)
        offsets = _arr(offsets, np.int32)
        n = len(offsets) - 1
        out = dict(level=np.empty(n, np.int32), url_id=np.empty(n, np.int32), sn=np.empty(n, np.int32),
                   time=np.empty(n, np.float64), flags=np.empty(n, np.uint8), status=np.empty(n, np.uint8))
        N.check(self._lib.hlsmap_view_from_json(self._h, n, _ptr(chars) if len(chars) else None, _ptr(offsets),
                                                _ptr(out["level"]), _ptr(out["url_id"]), _ptr(out["sn"]),
                                                _ptr(out["time"]), _ptr(out["flags"]), _ptr(out["status"])))
        return out

    # ---- media playlists in text form (hlsmap_parse_m3u8) ----------------
    @staticmethod
    def _text_args(text, text_off):
        if isinstance(text, (list, tuple)):          # one bytes object per track
            text_off = np.concatenate([[0], np.cumsum([len(t) for t in text])]).astype(np.int64)
            text = b"".join(text)
        buf = np.frombuffer(text, np.uint8) if isinstance(text, (bytes, bytearray, memoryview)) else _arr(text, np.uint8)
        return buf, _arr(text_off, np.int64)

    def parse_m3u8(self, text, text_off=None, track_state=None):
        """Level playlists (M3U8 text) -> SoA fragment columns, parsed on the GPU.  `text` is a list
        of bytes (one per track) or one buffer with `text_off`.  -> dict(track_off, live, start, sn,
        duration, result)"""
        buf, off = self._text_args(text, text_off)
        t = len(off) - 1
        st = None if track_state is None else _arr(track_state, np.uint8)
        track_off, live = np.zeros(t + 1, np.int64), np.ones(max(t, 1), np.uint8)
        res = N.IngestResult()
        args = [self._h, t, _ptr(st) if st is not None else None, _ptr(buf) if len(buf) else None, _ptr(off),
                _ptr(track_off), _ptr(live)]
        N.check(self._lib.hlsmap_parse_m3u8(*args, None, None, None, 0, ctypes.byref(res)))
        n = int(res.n_fragments)
        start, sn, dur = np.empty(n, np.float64), np.empty(n, np.int32), np.empty(n, np.float64)
        if n:
            N.check(self._lib.hlsmap_parse_m3u8(*args, _ptr(start), _ptr(sn), _ptr(dur), n, ctypes.byref(res)))
        return dict(track_off=track_off, live=live[:t], start=start, sn=sn, duration=dur, result=res)

    def upload_m3u8(self, playlist_base, track_state, text, text_off=None, url_count=None):
        """`new MediaMap(hls)` for playlists still in text form: parsed on the GPU, columns stay there."""
        return DeviceMap.from_m3u8(self, playlist_base, track_state, text, text_off, url_count)

    def pinned(self, shape, dtype):
        """numpy array over page-locked host memory (freed with the returned owner)."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        p = ctypes.c_void_p()
        N.check(self._lib.hlsmap_host_alloc(self._h, n, ctypes.byref(p)))
        buf = (ctypes.c_char * max(n, 1)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
        return _Pinned(self, p, arr)

    # ---- SegmentView / TrackView columns ------------------------------------------------
    def view_pack(self, level, url_id, sn):
        level, url_id, sn = _arr(level, np.int32), _arr(url_id, np.int32), _arr(sn, np.int32)
        wire = np.empty(3 * len(sn), np.uint32)
        N.check(self._lib.hlsmap_view_pack(self._h, len(sn), _ptr(level), _ptr(url_id), _ptr(sn),
                                           _ptr(wire)))
        return wire

    def view_unpack(self, wire):
        wire = _arr(wire, np.uint32)
        n = len(wire) // 3
        lv, ui, sn = np.empty(n, np.uint32), np.empty(n, np.uint32), np.empty(n, np.uint32)
        N.check(self._lib.hlsmap_view_unpack(self._h, n, _ptr(wire), _ptr(lv), _ptr(ui), _ptr(sn)))
        return lv, ui, sn

    def view_is_equal(self, al, au, as_, bl, bu, bs):
        arrs = [_arr(x, np.int32) for x in (al, au, as_, bl, bu, bs)]
        out = np.empty(len(arrs[0]), np.uint8)
        N.check(self._lib.hlsmap_view_is_equal(self._h, len(out), *[_ptr(a) for a in arrs], _ptr(out)))
        return out

    def view_is_in_track(self, level, url_id, t_level, t_url_id):
        arrs = [_arr(x, np.int32) for x in (level, url_id, t_level, t_url_id)]
        out = np.empty(len(arrs[0]), np.uint8)
        N.check(self._lib.hlsmap_view_is_in_track(self._h, len(out), *[_ptr(a) for a in arrs], _ptr(out)))
        return out

    def view_roundtrip(self, level, url_id, sn):
        level, url_id, sn = _arr(level, np.int32), _arr(url_id, np.int32), _arr(sn, np.int32)
        wire, eq = np.empty(3 * len(sn), np.uint32), np.empty(len(sn), np.uint8)
        N.check(self._lib.hlsmap_view_roundtrip(self._h, len(sn), _ptr(level), _ptr(url_id), _ptr(sn),
                                                _ptr(wire), _ptr(eq)))
        return wire, eq

    def view_format(self, level, url_id, sn=None, with_sn=True):
        level, url_id = _arr(level, np.int32), _arr(url_id, np.int32)
        sn = _arr(sn, np.int32) if with_sn else None
        n = len(level)
        offsets = np.empty(n + 1, np.int32)
        cap = 36 * n
        chars = np.empty(max(cap, 1), np.uint8)
        total = ctypes.c_int64(0)
        N.check(self._lib.hlsmap_view_format(self._h, n, _ptr(level), _ptr(url_id), _ptr(sn),
                                             int(bool(with_sn)), _ptr(offsets), _ptr(chars), cap,
                                             ctypes.byref(total)))
        return offsets, chars[:total.value].copy()


class _Pinned:
    def __init__(self, ctx, ptr, arr):
        self.ctx, self.ptr, self.array = ctx, ptr, arr

    def free(self):
        if self.ptr is not None and self.ctx._h:
            self.array = None
            N.check(self.ctx._lib.hlsmap_host_free(self.ctx._h, self.ptr))
        self.ptr = None


class DeviceMap:
    """Fragment tables resident in HBM: the batch analogue of `new MediaMap(hls)`."""

    def __init__(self, ctx: Context, tables: FragmentTables):
        tables.validate()
        self.ctx = ctx
        self._lib = ctx._lib
        self.n_playlists = tables.n_playlists
        h = ctypes.c_void_p()
        t = tables
        pb = _arr(t.playlist_base, np.int32)
        to = _arr(t.track_off, np.int64)
        ts = _arr(t.track_state, np.uint8)
        uc = _arr(t.url_count, np.int32)
        st, sn, du = _arr(t.start, np.float64), _arr(t.sn, np.int32), _arr(t.duration, np.float64)
        N.check(self._lib.hlsmap_map_create(ctx._h, t.n_playlists, _ptr(pb), t.n_tracks, _ptr(to),
                                            _ptr(ts), _ptr(uc), _ptr(st), _ptr(sn), _ptr(du),
                                            ctypes.byref(h)))
        self._h = h

    @classmethod
    def from_m3u8(cls, ctx, playlist_base, track_state, text, text_off=None, url_count=None):
        buf, off = Context._text_args(text, text_off)
        pb, ts = _arr(playlist_base, np.int32), _arr(track_state, np.uint8)
        uc = None if url_count is None else _arr(url_count, np.int32)
        self = object.__new__(cls)
        self.ctx, self._lib, self.n_playlists = ctx, ctx._lib, len(pb) - 1
        self.ingest = N.IngestResult()
        h = ctypes.c_void_p()
        N.check(self._lib.hlsmap_map_create_from_m3u8(
            ctx._h, len(pb) - 1, _ptr(pb), len(ts), _ptr(ts), _ptr(uc) if uc is not None else None,
            _ptr(buf) if len(buf) else None, _ptr(off), ctypes.byref(self.ingest), ctypes.byref(h)))
        self._h = h
        return self

    def close(self):
        if getattr(self, "_h", None) and self.ctx._h:
            self._lib.hlsmap_map_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        info = N.MapInfo()
        N.check(self._lib.hlsmap_map_get_info(self._h, ctypes.byref(info)))
        return {k: getattr(info, k) for k, _ in N.MapInfo._fields_}

    def monotone_flags(self):
        out = np.zeros(self.info()["n_tracks"], np.uint8)
        N.check(self._lib.hlsmap_map_get_monotone(self._h, _ptr(out)))
        return out

    def set_path(self, path):
        N.check(self._lib.hlsmap_map_set_path(self._h, int(path)))

    def slide_tracks(self, playlist, level, drop, append_offset, new_start, new_sn, new_dur0=None):
        """Live update: per listed track evict `drop` fragments at the front and append the
        fragments new_start/new_sn[append_offset[i]:append_offset[i+1]] (media-map.js:28,41 reads
        hls.levels live; this keeps the HBM copy current without re-uploading the map)."""
        playlist, level, drop = _arr(playlist, np.int32), _arr(level, np.int32), _arr(drop, np.int32)
        append_offset = _arr(append_offset, np.int64)
        new_start, new_sn = _arr(new_start, np.float64), _arr(new_sn, np.int32)
        new_dur0 = _arr(new_dur0, np.float64)
        N.check(self._lib.hlsmap_map_slide_tracks(self._h, len(level), _ptr(playlist), _ptr(level),
                                                  _ptr(drop), _ptr(append_offset), _ptr(new_start),
                                                  _ptr(new_sn), _ptr(new_dur0)))

    # ---- getSegmentList -------------------------------------------------------------------
    def get_segment_list(self, playlist, level, begin, duration, want_status=True,
                         capacity=None):
        """Host arrays in, host arrays out (list form).  Grows the hit buffers and retries
        once if the first capacity guess was too small."""
        playlist = _arr(playlist, np.int32)
        level, begin, duration = _arr(level, np.int32), _arr(begin, np.float64), _arr(duration, np.float64)
        nq = len(level)
        offsets = np.empty(nq + 1, np.int32)
        status = np.empty(nq, np.uint8) if want_status else None
        cap = int(capacity) if capacity is not None else 16 * nq + 1024
        res = N.Result()
        for _ in range(2):
            sn = np.empty(max(cap, 1), np.int32)
            tm = np.empty(max(cap, 1), np.float64)
            code = self._lib.hlsmap_get_segment_list(
                self._h, nq, _ptr(playlist), _ptr(level), _ptr(begin), _ptr(duration), _ptr(offsets),
                _ptr(sn), _ptr(tm), cap, _ptr(status), ctypes.byref(res))
            if code == N.ERR_CAPACITY and capacity is None:
                cap = int(res.total_hits)
                continue
            N.check(code)
            break
        out = SegmentListResult()
        out.offsets, out.status = offsets, status
        out.sn, out.time = sn[:res.total_hits], tm[:res.total_hits]
        out.total_hits, out.n_unparsed, out.n_missing = res.total_hits, res.n_unparsed, res.n_missing
        out.kernel_launches, out.path = res.kernel_launches, res.path
        return out

    def get_segment_list_into(self, playlist, level, begin, duration, offsets, sn, time,
                              status=None):
        """The raw host-pointer call on caller-owned (ideally pinned) numpy arrays: the
        end-to-end path a binding takes.  Returns the Result; raises on capacity errors."""
        res = N.Result()
        N.check(self._lib.hlsmap_get_segment_list(
            self._h, len(level), _ptr(playlist), _ptr(level), _ptr(begin), _ptr(duration),
            _ptr(offsets), _ptr(sn), _ptr(time), len(sn), _ptr(status), ctypes.byref(res)))
        return res

    def count_segments(self, playlist, level, begin, duration):
        """Count-only pass: offsets[Q+1] and the totals, no hit arrays."""
        playlist = _arr(playlist, np.int32)
        level, begin, duration = _arr(level, np.int32), _arr(begin, np.float64), _arr(duration, np.float64)
        nq = len(level)
        offsets = np.empty(nq + 1, np.int32)
        res = N.Result()
        N.check(self._lib.hlsmap_get_segment_list(
            self._h, nq, _ptr(playlist), _ptr(level), _ptr(begin), _ptr(duration), _ptr(offsets),
            None, None, 0, None, ctypes.byref(res)))
        return offsets, res.total_hits

    def get_segment_list_dev(self, nq, d_playlist, d_level, d_begin, d_duration, d_offsets,
                             d_sn_out, d_time_out, capacity, d_status=0, stream=None, sync=False):
        """Device pointers (ints); asynchronous unless sync=True (then returns the Result)."""
        res = N.Result() if sync else None
        code = self._lib.hlsmap_get_segment_list_dev(
            self._h, int(nq), d_playlist or None, d_level, d_begin, d_duration, d_offsets,
            d_sn_out or None, d_time_out or None, int(capacity), d_status or None, _stream(stream),
            ctypes.byref(res) if sync else None)
        N.check(code)
        return res

    def get_segment_list_dev_batches(self, batches, stream=None, event_begin=None, event_end=None):
        """Several independent device-pointer batches with one call into the library.  `batches`: a
        sequence of dicts with nq, d_playlist (or 0), d_level, d_begin, d_duration, d_offsets, d_sn_out,
        d_time_out, capacity, d_status (optional) -- or a ready-made ctypes array of N.Batch.  Asynchronous;
        returns the number of batches launched.  event_begin / event_end: raw cudaEvent_t handles (ints)
        recorded on the stream right before the first and right after the last launch, inside the call."""
        if not isinstance(batches, ctypes.Array):
            arr = (N.Batch * len(batches))()
            for a, b in zip(arr, batches):
                a.nq, a.d_playlist, a.d_level = int(b["nq"]), b.get("d_playlist") or None, b["d_level"]
                a.d_begin, a.d_duration, a.d_offsets = b["d_begin"], b["d_duration"], b["d_offsets"]
                a.d_sn_out, a.d_time_out = b.get("d_sn_out") or None, b.get("d_time_out") or None
                a.hit_capacity, a.d_status = int(b.get("capacity", 0)), b.get("d_status") or None
            batches = arr
        done = ctypes.c_int32(0)
        N.check(self._lib.hlsmap_get_segment_list_dev_batches_timed(
            self._h, len(batches), batches, _stream(stream),
            ctypes.c_void_p(event_begin) if event_begin else None,
            ctypes.c_void_p(event_end) if event_end else None, ctypes.byref(done)))
        return done.value

    def get_segment_list_noalloc_dev(self, nq, d_playlist, d_level, d_begin, d_duration, d_offsets,
                                     d_sn_out, d_time_out, capacity, d_status=0, stream=None,
                                     sync=True):
        """Same batch through the entry point that never allocates (for hosts that forbid
        cudaMalloc while running): fixed in-library scratch, at most 65536 tiles per call, calls
        must not overlap on the GPU.  Device pointers as ints."""
        res = N.Result() if sync else None
        N.check(self._lib.hlsmap_get_segment_list_noalloc_dev(
            self._h, int(nq), d_playlist or None, d_level, d_begin, d_duration, d_offsets,
            d_sn_out or None, d_time_out or None, int(capacity), d_status or None, _stream(stream),
            ctypes.byref(res) if sync else None))
        return res

    def last_result(self, stream=None):
        res = N.Result()
        N.check(self._lib.hlsmap_ctx_last_result(self.ctx._h, _stream(stream), ctypes.byref(res)))
        return res

    def get_segment_range(self, playlist, level, begin, duration):
        playlist = _arr(playlist, np.int32)
        level, begin, duration = _arr(level, np.int32), _arr(begin, np.float64), _arr(duration, np.float64)
        nq = len(level)
        first, count, status = np.empty(nq, np.int32), np.empty(nq, np.int32), np.empty(nq, np.uint8)
        N.check(self._lib.hlsmap_get_segment_range(self._h, nq, _ptr(playlist), _ptr(level), _ptr(begin),
                                                   _ptr(duration), _ptr(first), _ptr(count), _ptr(status)))
        return first, count, status

    def get_segment_duration(self, playlist, level):
        playlist, level = _arr(playlist, np.int32), _arr(level, np.int32)
        n = len(level)
        out, status = np.empty(n, np.float64), np.empty(n, np.uint8)
        N.check(self._lib.hlsmap_get_segment_duration(self._h, n, _ptr(playlist), _ptr(level),
                                                      _ptr(out), _ptr(status)))
        return out, status

    def views_from_fragments(self, playlist, level, index):
        """(sn, time, status) of fragment `index` of each (playlist, level): what the loader puts
        into `new SegmentView({sn: frag.sn, trackView, time: frag.start})`."""
        playlist, level, index = _arr(playlist, np.int32), _arr(level, np.int32), _arr(index, np.int32)
        n = len(level)
        sn, tm, st = np.empty(n, np.int32), np.empty(n, np.float64), np.empty(n, np.uint8)
        N.check(self._lib.hlsmap_views_from_fragments(self._h, n, _ptr(playlist), _ptr(level), _ptr(index),
                                                      _ptr(sn), _ptr(tm), _ptr(st)))
        return sn, tm, st

    def get_next_segment_view(self, playlist, level, sn):
        """EXTENSION (not in the reference, see include/hlsmap.h): the fragment after the first
        fragment whose sn matches -> (sn, time, found, status)."""
        playlist, level, sn = _arr(playlist, np.int32), _arr(level, np.int32), _arr(sn, np.int32)
        n = len(level)
        osn, tm = np.empty(n, np.int32), np.empty(n, np.float64)
        fd, st = np.empty(n, np.uint8), np.empty(n, np.uint8)
        N.check(self._lib.hlsmap_get_next_segment_view(self._h, n, _ptr(playlist), _ptr(level), _ptr(sn),
                                                       _ptr(osn), _ptr(tm), _ptr(fd), _ptr(st)))
        return osn, tm, fd, st

    def get_track_list(self, playlist=0):
        n = ctypes.c_int64(0)
        N.check(self._lib.hlsmap_get_track_list(self._h, int(playlist), None, None, 0, ctypes.byref(n)))
        lv, ui = np.empty(n.value, np.int32), np.empty(n.value, np.int32)
        if n.value:
            N.check(self._lib.hlsmap_get_track_list(self._h, int(playlist), _ptr(lv), _ptr(ui),
                                                    n.value, ctypes.byref(n)))
        return lv, ui

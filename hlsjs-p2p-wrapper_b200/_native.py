"""ctypes binding of csrc/libhlsmap.so -- the same C ABI (include/hlsmap.h) an N-API addon
binds.  There is no fallback: if the library is missing or no CUDA device is usable,
everything here raises."""
import ctypes
import os

from . import build as _build

c_i32p = ctypes.c_void_p
OK = 0
ERR_INVALID, ERR_CUDA, ERR_NOMEM, ERR_CAPACITY, ERR_OVERFLOW, ERR_UNSUPPORTED = -1, -2, -3, -4, -5, -6
ST_OK, ST_UNPARSED, ST_MISSING, ST_EMPTY, ST_SCATTERED, ST_NO_FRAGMENT = 0, 1, 2, 3, 4, 5
TRACK_MISSING, TRACK_UNPARSED, TRACK_PARSED = 0, 1, 2
PATH_AUTO, PATH_SEARCH, PATH_FILTER = 0, 1, 2
JSON_OK, JSON_MALFORMED, JSON_UNSUPPORTED, JSON_NOT_INT32 = 0, 1, 2, 3
JSON_HAS_SN, JSON_HAS_TRACKVIEW, JSON_HAS_LEVEL, JSON_HAS_URLID, JSON_HAS_TIME, JSON_TIME_NULL = 1, 2, 4, 8, 16, 32


class Result(ctypes.Structure):
    _fields_ = [("total_hits", ctypes.c_int64), ("n_unparsed", ctypes.c_int64),
                ("n_missing", ctypes.c_int64), ("kernel_launches", ctypes.c_int32),
                ("path", ctypes.c_int32)]


class MapInfo(ctypes.Structure):
    _fields_ = [("n_playlists", ctypes.c_int32), ("n_tracks", ctypes.c_int32),
                ("n_fragments", ctypes.c_int64), ("n_monotone_tracks", ctypes.c_int32),
                ("n_parsed_tracks", ctypes.c_int32), ("all_monotone", ctypes.c_int32),
                ("stage_all", ctypes.c_int32), ("device_bytes", ctypes.c_int64)]


class Batch(ctypes.Structure):
    """hlsmap_batch: one list-form batch of hlsmap_get_segment_list_dev_batches (device pointers)."""
    _fields_ = [("nq", ctypes.c_int64), ("d_playlist", ctypes.c_void_p), ("d_level", ctypes.c_void_p),
                ("d_begin", ctypes.c_void_p), ("d_duration", ctypes.c_void_p), ("d_offsets", ctypes.c_void_p),
                ("d_sn_out", ctypes.c_void_p), ("d_time_out", ctypes.c_void_p), ("hit_capacity", ctypes.c_int64),
                ("d_status", ctypes.c_void_p)]


class IngestResult(ctypes.Structure):
    _fields_ = [("n_fragments", ctypes.c_int64), ("n_sequence_tags", ctypes.c_int64),
                ("n_unsupported", ctypes.c_int64), ("text_bytes", ctypes.c_int64),
                ("kernel_launches", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class HlsMapError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("libhlsmap error %d: %s" % (code, message))
        self.code = code
        self.message = message


# every symbol include/hlsmap.h declares: name -> (restype, argtypes)
_P, _I64, _I32, _INT = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_int
SYMBOLS = {
    "hlsmap_last_error": (ctypes.c_char_p, []),
    "hlsmap_version": (_INT, []),
    "hlsmap_device_count": (_INT, []),
    "hlsmap_ctx_create": (_INT, [_INT, _P]),
    "hlsmap_ctx_destroy": (None, [_P]),
    "hlsmap_ctx_device": (_INT, [_P]),
    "hlsmap_ctx_set_batch_overlap": (_INT, [_P, _INT]),
    "hlsmap_ctx_synchronize": (_INT, [_P]),
    "hlsmap_host_alloc": (_INT, [_P, _I64, _P]),
    "hlsmap_host_free": (_INT, [_P, _P]),
    "hlsmap_map_create": (_INT, [_P, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P]),
    "hlsmap_map_destroy": (None, [_P]),
    "hlsmap_map_get_info": (_INT, [_P, _P]),
    "hlsmap_map_get_monotone": (_INT, [_P, _P]),
    "hlsmap_map_set_path": (_INT, [_P, _INT]),
    "hlsmap_map_slide_tracks": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P]),
    "hlsmap_get_segment_list_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P, _I64, _P, _P, _P]),
    "hlsmap_get_segment_list": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P, _I64, _P, _P]),
    "hlsmap_get_segment_list_noalloc_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P, _I64, _P, _P, _P]),
    "hlsmap_get_segment_list_dev_batches": (_INT, [_P, _I32, _P, _P, _P]),
    "hlsmap_get_segment_list_dev_batches_timed": (_INT, [_P, _I32, _P, _P, _P, _P, _P]),
    "hlsmap_ctx_last_result": (_INT, [_P, _P, _P]),
    "hlsmap_get_segment_range_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P, _P]),
    "hlsmap_get_segment_range": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P]),
    "hlsmap_get_segment_duration_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P]),
    "hlsmap_get_segment_duration": (_INT, [_P, _I64, _P, _P, _P, _P]),
    "hlsmap_views_from_fragments_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P]),
    "hlsmap_views_from_fragments": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P]),
    "hlsmap_get_next_segment_view_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P, _P]),
    "hlsmap_get_next_segment_view": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P]),
    "hlsmap_view_to_json_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _I64, _P, _P]),
    "hlsmap_view_to_json": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _I64, _P]),
    "hlsmap_view_from_json_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "hlsmap_view_from_json": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P, _P]),
    "hlsmap_parse_m3u8_dev": (_INT, [_P, _I32, _P, _P, _P, _I64, _P, _P, _P, _P, _P, _I64, _P, _P]),
    "hlsmap_parse_m3u8": (_INT, [_P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _P]),
    "hlsmap_map_create_from_m3u8": (_INT, [_P, _I32, _P, _I32, _P, _P, _P, _P, _P, _P]),
    "hlsmap_map_create_from_m3u8_dev": (_INT, [_P, _I32, _P, _I32, _P, _P, _P, _P, _I64, _P, _P]),
    "hlsmap_get_track_list": (_INT, [_P, _I32, _P, _P, _I64, _P]),
    "hlsmap_view_pack_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P]),
    "hlsmap_view_pack": (_INT, [_P, _I64, _P, _P, _P, _P]),
    "hlsmap_view_unpack_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P]),
    "hlsmap_view_unpack": (_INT, [_P, _I64, _P, _P, _P, _P]),
    "hlsmap_view_is_equal_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P, _P]),
    "hlsmap_view_is_equal": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P, _P]),
    "hlsmap_view_is_in_track_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P]),
    "hlsmap_view_is_in_track": (_INT, [_P, _I64, _P, _P, _P, _P, _P]),
    "hlsmap_view_roundtrip_dev": (_INT, [_P, _I64, _P, _P, _P, _P, _P, _P]),
    "hlsmap_view_roundtrip": (_INT, [_P, _I64, _P, _P, _P, _P, _P]),
    "hlsmap_view_format_dev": (_INT, [_P, _I64, _P, _P, _P, _INT, _P, _P, _I64, _P, _P]),
    "hlsmap_view_format": (_INT, [_P, _I64, _P, _P, _P, _INT, _P, _P, _I64, _P]),
}

_lib = None


def library_path():
    return _build.LIB


def load():
    """dlopen libhlsmap.so and type every entry point.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        raise ImportError(
            "%s is missing: build it first (python -c 'import __graft_entry__ as g; g.build()'). "
            "This package has no CPU fallback." % path)
    lib = ctypes.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the ABI lost a symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(code):
    """Raise HlsMapError for a negative status; returns the code otherwise."""
    if code < 0:
        msg = load().hlsmap_last_error()
        raise HlsMapError(code, msg.decode("utf-8", "replace") if msg else "")
    return code

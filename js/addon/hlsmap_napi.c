/*
 * hlsmap_napi.c -- N-API addon: the thin JavaScript <-> C-ABI layer of libhlsmap.
 *
 * Zero logic lives here: every function unwraps typed arrays, calls one entry point of
 * include/hlsmap.h and wraps the outputs in typed arrays.  js/mapping.js builds the
 * reference's classes (MediaMap / SegmentView / TrackView) on top.
 *
 * Build where Node exists (not in this image: no node, no node_api.h):
 *     cd js/addon && node-gyp configure build      (binding.gyp links ../../hlsjs-p2p-wrapper_b200/csrc/libhlsmap.so)
 * `gcc -fsyntax-only -I. -I../../include hlsmap_napi.c` checks it here against the minimal
 * declarations in node_api_min.h.
 */
#if defined(__has_include)
#if __has_include(<node_api.h>)
#include <node_api.h>
#else
#include "node_api_min.h"
#endif
#else
#include <node_api.h>
#endif
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "hlsmap.h"

#define CHECK(env, call)                                                  \
    do {                                                                  \
        if ((call) != napi_ok) {                                          \
            napi_throw_error((env), NULL, "hlsmap addon: N-API call failed: " #call); \
            return NULL;                                                  \
        }                                                                 \
    } while (0)

static napi_value fail_hlsmap(napi_env env)
{
    napi_throw_error(env, "HLSMAP", hlsmap_last_error());
    return NULL;
}

/* typed array (or null/undefined -> NULL pointer) -> data pointer + element count */
static int typed(napi_env env, napi_value v, napi_typedarray_type want, void **data, size_t *len)
{
    napi_valuetype t;
    *data = NULL;
    *len = 0;
    if (napi_typeof(env, v, &t) != napi_ok) return 0;
    if (t == napi_null || t == napi_undefined) return 1;
    bool is = false;
    if (napi_is_typedarray(env, v, &is) != napi_ok || !is) return 0;
    napi_typedarray_type got;
    napi_value ab;
    size_t off;
    if (napi_get_typedarray_info(env, v, &got, len, data, &ab, &off) != napi_ok) return 0;
    return got == want;
}

static napi_value new_typed(napi_env env, napi_typedarray_type type, size_t elem, size_t n, void **data)
{
    napi_value ab, ta;
    if (napi_create_arraybuffer(env, n * elem, data, &ab) != napi_ok) return NULL;
    if (napi_create_typedarray(env, type, n, ab, 0, &ta) != napi_ok) return NULL;
    return ta;
}

static napi_value field(napi_env env, napi_value obj, const char *name)
{
    napi_value v = NULL;
    napi_get_named_property(env, obj, name, &v);
    return v;
}

static void free_ctx(napi_env env, void *p, void *hint) { (void)env; (void)hint; hlsmap_ctx_destroy((hlsmap_ctx *)p); }

/* createContext(device) -> external */
static napi_value CreateContext(napi_env env, napi_callback_info info)
{
    size_t argc = 1;
    napi_value argv[1], out;
    int32_t device = 0;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    if (argc > 0) napi_get_value_int32(env, argv[0], &device);
    hlsmap_ctx *ctx = NULL;
    if (hlsmap_ctx_create(device, &ctx) != HLSMAP_OK) return fail_hlsmap(env);
    CHECK(env, napi_create_external(env, ctx, free_ctx, NULL, &out));
    return out;
}

/* createMap(ctx, {playlistBase, trackOffset, trackState, urlCount, start, sn, duration}) -> external
 * (`new MediaMap(hls)`, media-map.js:6-8) */
static napi_value CreateMap(napi_env env, napi_callback_info info)
{
    size_t argc = 2, n;
    napi_value argv[2], out;
    void *ctx, *pb, *to, *ts, *uc, *st, *sn, *du;
    size_t npb, nto, nts, nuc, nst, nsn, ndu;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    if (!typed(env, field(env, argv[1], "playlistBase"), napi_int32_array, &pb, &npb) ||
        !typed(env, field(env, argv[1], "trackOffset"), napi_bigint64_array, &to, &nto) ||
        !typed(env, field(env, argv[1], "trackState"), napi_uint8_array, &ts, &nts) ||
        !typed(env, field(env, argv[1], "urlCount"), napi_int32_array, &uc, &nuc) ||
        !typed(env, field(env, argv[1], "start"), napi_float64_array, &st, &nst) ||
        !typed(env, field(env, argv[1], "sn"), napi_int32_array, &sn, &nsn) ||
        !typed(env, field(env, argv[1], "duration"), napi_float64_array, &du, &ndu) || npb < 1 || nto < 1) {
        napi_throw_type_error(env, NULL, "createMap: expected typed-array tables");
        return NULL;
    }
    (void)n;
    hlsmap_map *map = NULL;
    if (hlsmap_map_create((hlsmap_ctx *)ctx, (int32_t)(npb - 1), (const int32_t *)pb, (int32_t)(nto - 1),
                          (const int64_t *)to, (const uint8_t *)ts, (const int32_t *)uc,
                          (const double *)st, (const int32_t *)sn, (const double *)du, &map) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_external(env, map, NULL, NULL, &out));
    return out;
}

static napi_value DestroyMap(napi_env env, napi_callback_info info)
{
    size_t argc = 1;
    napi_value argv[1];
    void *map;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &map));
    hlsmap_map_destroy((hlsmap_map *)map);
    return NULL;
}

/* getSegmentList(map, playlist|null, level, begin, duration)
 *   -> {offsets, sn, time, status, totalHits, nUnparsed, nMissing}   (media-map.js:27-54, batched) */
static napi_value GetSegmentList(napi_env env, napi_callback_info info)
{
    size_t argc = 5, nq, n;
    napi_value argv[5], out, v;
    void *map, *pl, *lv, *bg, *du, *off, *st, *sn, *tm;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &map));
    if (!typed(env, argv[1], napi_int32_array, &pl, &n) || !typed(env, argv[2], napi_int32_array, &lv, &nq) ||
        !typed(env, argv[3], napi_float64_array, &bg, &n) || n != nq ||
        !typed(env, argv[4], napi_float64_array, &du, &n) || n != nq) {
        napi_throw_type_error(env, NULL, "getSegmentList: expected Int32Array level, Float64Array begin/duration");
        return NULL;
    }
    napi_value a_off = new_typed(env, napi_int32_array, 4, nq + 1, &off);
    napi_value a_st = new_typed(env, napi_uint8_array, 1, nq, &st);
    if (!a_off || !a_st) return NULL;
    hlsmap_result res;
    /* two-call pattern: count, then fill buffers of exactly the needed size */
    if (hlsmap_get_segment_list((hlsmap_map *)map, (int64_t)nq, (const int32_t *)pl, (const int32_t *)lv,
                                (const double *)bg, (const double *)du, (int32_t *)off, NULL, NULL, 0,
                                (uint8_t *)st, &res) != HLSMAP_OK)
        return fail_hlsmap(env);
    napi_value a_sn = new_typed(env, napi_int32_array, 4, (size_t)res.total_hits, &sn);
    napi_value a_tm = new_typed(env, napi_float64_array, 8, (size_t)res.total_hits, &tm);
    if (!a_sn || !a_tm) return NULL;
    if (res.total_hits > 0 &&
        hlsmap_get_segment_list((hlsmap_map *)map, (int64_t)nq, (const int32_t *)pl, (const int32_t *)lv,
                                (const double *)bg, (const double *)du, (int32_t *)off, (int32_t *)sn,
                                (double *)tm, res.total_hits, (uint8_t *)st, &res) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_object(env, &out));
    napi_set_named_property(env, out, "offsets", a_off);
    napi_set_named_property(env, out, "sn", a_sn);
    napi_set_named_property(env, out, "time", a_tm);
    napi_set_named_property(env, out, "status", a_st);
    napi_create_double(env, (double)res.total_hits, &v);
    napi_set_named_property(env, out, "totalHits", v);
    napi_create_double(env, (double)res.n_unparsed, &v);
    napi_set_named_property(env, out, "nUnparsed", v);
    napi_create_double(env, (double)res.n_missing, &v);
    napi_set_named_property(env, out, "nMissing", v);
    return out;
}

/* getSegmentDuration(map, playlist|null, level) -> {out, status}   (media-map.js:81-87) */
static napi_value GetSegmentDuration(napi_env env, napi_callback_info info)
{
    size_t argc = 3, nq, n;
    napi_value argv[3], out;
    void *map, *pl, *lv, *o, *st;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &map));
    if (!typed(env, argv[1], napi_int32_array, &pl, &n) || !typed(env, argv[2], napi_int32_array, &lv, &nq)) {
        napi_throw_type_error(env, NULL, "getSegmentDuration: expected Int32Array level");
        return NULL;
    }
    napi_value a_o = new_typed(env, napi_float64_array, 8, nq, &o);
    napi_value a_st = new_typed(env, napi_uint8_array, 1, nq, &st);
    if (!a_o || !a_st) return NULL;
    if (hlsmap_get_segment_duration((hlsmap_map *)map, (int64_t)nq, (const int32_t *)pl, (const int32_t *)lv,
                                    (double *)o, (uint8_t *)st) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_object(env, &out));
    napi_set_named_property(env, out, "out", a_o);
    napi_set_named_property(env, out, "status", a_st);
    return out;
}

/* getTrackList(map, playlist) -> {level, urlId}   (media-map.js:60-73) */
static napi_value GetTrackList(napi_env env, napi_callback_info info)
{
    size_t argc = 2;
    napi_value argv[2], out;
    void *map, *lv, *ui;
    int32_t playlist = 0;
    int64_t n = 0;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &map));
    if (argc > 1) napi_get_value_int32(env, argv[1], &playlist);
    if (hlsmap_get_track_list((hlsmap_map *)map, playlist, NULL, NULL, 0, &n) != HLSMAP_OK) return fail_hlsmap(env);
    napi_value a_lv = new_typed(env, napi_int32_array, 4, (size_t)n, &lv);
    napi_value a_ui = new_typed(env, napi_int32_array, 4, (size_t)n, &ui);
    if (!a_lv || !a_ui) return NULL;
    if (n > 0 && hlsmap_get_track_list((hlsmap_map *)map, playlist, (int32_t *)lv, (int32_t *)ui, n, &n) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_object(env, &out));
    napi_set_named_property(env, out, "level", a_lv);
    napi_set_named_property(env, out, "urlId", a_ui);
    return out;
}

/* viewPack(ctx, level, urlId, sn) -> Uint32Array(3n)   (segment-view.js:59-61, batched) */
static napi_value ViewPack(napi_env env, napi_callback_info info)
{
    size_t argc = 4, n, m;
    napi_value argv[4];
    void *ctx, *lv, *ui, *sn, *wire;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    if (!typed(env, argv[1], napi_int32_array, &lv, &n) || !typed(env, argv[2], napi_int32_array, &ui, &m) ||
        m != n || !typed(env, argv[3], napi_int32_array, &sn, &m) || m != n) {
        napi_throw_type_error(env, NULL, "viewPack: expected three Int32Array of equal length");
        return NULL;
    }
    napi_value a = new_typed(env, napi_uint32_array, 4, 3 * n, &wire);
    if (!a) return NULL;
    if (hlsmap_view_pack((hlsmap_ctx *)ctx, (int64_t)n, (const int32_t *)lv, (const int32_t *)ui,
                         (const int32_t *)sn, (uint32_t *)wire) != HLSMAP_OK)
        return fail_hlsmap(env);
    return a;
}

/* viewUnpack(ctx, wire) -> {level, urlId, sn} as Uint32Array   (segment-view.js:9-17, batched) */
static napi_value ViewUnpack(napi_env env, napi_callback_info info)
{
    size_t argc = 2, n;
    napi_value argv[2], out;
    void *ctx, *wire, *lv, *ui, *sn;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    if (!typed(env, argv[1], napi_uint32_array, &wire, &n)) {
        napi_throw_type_error(env, NULL, "viewUnpack: expected a Uint32Array");
        return NULL;
    }
    n /= 3;
    napi_value a_lv = new_typed(env, napi_uint32_array, 4, n, &lv);
    napi_value a_ui = new_typed(env, napi_uint32_array, 4, n, &ui);
    napi_value a_sn = new_typed(env, napi_uint32_array, 4, n, &sn);
    if (!a_lv || !a_ui || !a_sn) return NULL;
    if (hlsmap_view_unpack((hlsmap_ctx *)ctx, (int64_t)n, (const uint32_t *)wire, (uint32_t *)lv,
                           (uint32_t *)ui, (uint32_t *)sn) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_object(env, &out));
    napi_set_named_property(env, out, "level", a_lv);
    napi_set_named_property(env, out, "urlId", a_ui);
    napi_set_named_property(env, out, "sn", a_sn);
    return out;
}

/* viewIsEqual(ctx, aLevel, aUrlId, aSn, bLevel, bUrlId, bSn) -> Uint8Array   (segment-view.js:33-39) */
static napi_value ViewIsEqual(napi_env env, napi_callback_info info)
{
    size_t argc = 7, n = 0, m;
    napi_value argv[7];
    void *ctx, *p[6], *o;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    for (int i = 0; i < 6; i++) {
        if (!typed(env, argv[1 + i], napi_int32_array, &p[i], &m) || (i && m != n)) {
            napi_throw_type_error(env, NULL, "viewIsEqual: expected six Int32Array of equal length");
            return NULL;
        }
        n = m;
    }
    napi_value a = new_typed(env, napi_uint8_array, 1, n, &o);
    if (!a) return NULL;
    if (hlsmap_view_is_equal((hlsmap_ctx *)ctx, (int64_t)n, (const int32_t *)p[0], (const int32_t *)p[1],
                             (const int32_t *)p[2], (const int32_t *)p[3], (const int32_t *)p[4],
                             (const int32_t *)p[5], (uint8_t *)o) != HLSMAP_OK)
        return fail_hlsmap(env);
    return a;
}

/* viewFormat(ctx, level, urlId, sn, withSn) -> {offsets: Int32Array(n+1), chars: Uint8Array}
 * (segment-view.js:52-54 / track-view.js:11-13, batched) */
static napi_value ViewFormat(napi_env env, napi_callback_info info)
{
    size_t argc = 5, n, m;
    napi_value argv[5], out;
    void *ctx, *lv, *ui, *sn, *off, *chars;
    bool with_sn = true;
    int64_t total = 0;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    if (!typed(env, argv[1], napi_int32_array, &lv, &n) || !typed(env, argv[2], napi_int32_array, &ui, &m) ||
        m != n || !typed(env, argv[3], napi_int32_array, &sn, &m)) {
        napi_throw_type_error(env, NULL, "viewFormat: expected Int32Array columns");
        return NULL;
    }
    if (argc > 4) napi_get_value_bool(env, argv[4], &with_sn);
    napi_value a_off = new_typed(env, napi_int32_array, 4, n + 1, &off);
    napi_value a_ch = new_typed(env, napi_uint8_array, 1, 36 * n + 1, &chars);
    if (!a_off || !a_ch) return NULL;
    if (hlsmap_view_format((hlsmap_ctx *)ctx, (int64_t)n, (const int32_t *)lv, (const int32_t *)ui,
                           (const int32_t *)sn, with_sn ? 1 : 0, (int32_t *)off, (char *)chars,
                           (int64_t)(36 * n + 1), &total) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_object(env, &out));
    napi_set_named_property(env, out, "offsets", a_off);
    napi_set_named_property(env, out, "chars", a_ch);
    return out;
}

/* createMapFromM3u8(ctx, {playlistBase, trackState, urlCount|null, text: Uint8Array, textOffset: BigInt64Array})
 *   -> external.  `new MediaMap(hls)` for playlists still in text form: the bytes hls.js 0.5.46 would
 *   have parsed (package.json:34) go to the GPU and the fragment tables are built there. */
static napi_value CreateMapFromM3u8(napi_env env, napi_callback_info info)
{
    size_t argc = 2, npb, nts, nuc, ntx, nto;
    napi_value argv[2], out;
    void *ctx, *pb, *ts, *uc, *tx, *to;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    if (!typed(env, field(env, argv[1], "playlistBase"), napi_int32_array, &pb, &npb) ||
        !typed(env, field(env, argv[1], "trackState"), napi_uint8_array, &ts, &nts) ||
        !typed(env, field(env, argv[1], "urlCount"), napi_int32_array, &uc, &nuc) ||
        !typed(env, field(env, argv[1], "text"), napi_uint8_array, &tx, &ntx) ||
        !typed(env, field(env, argv[1], "textOffset"), napi_bigint64_array, &to, &nto) || npb < 1 || nto != nts + 1) {
        napi_throw_type_error(env, NULL, "createMapFromM3u8: expected typed-array inputs");
        return NULL;
    }
    hlsmap_map *map = NULL;
    hlsmap_ingest_result res;
    if (hlsmap_map_create_from_m3u8((hlsmap_ctx *)ctx, (int32_t)(npb - 1), (const int32_t *)pb, (int32_t)nts,
                                    (const uint8_t *)ts, (const int32_t *)uc, (const char *)tx, (const int64_t *)to,
                                    &res, &map) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_external(env, map, NULL, NULL, &out));
    return out;
}

/* viewsFromFragments(map, playlist|null, level, index) -> {sn, time, status}
 * (`new SegmentView({sn: frag.sn, trackView, time: frag.start})`, p2p-loader-generator.js:145-147, batched) */
static napi_value ViewsFromFragments(napi_env env, napi_callback_info info)
{
    size_t argc = 4, n, m;
    napi_value argv[4], out;
    void *map, *pl, *lv, *ix, *sn, *tm, *st;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &map));
    if (!typed(env, argv[1], napi_int32_array, &pl, &m) || !typed(env, argv[2], napi_int32_array, &lv, &n) ||
        !typed(env, argv[3], napi_int32_array, &ix, &m) || m != n) {
        napi_throw_type_error(env, NULL, "viewsFromFragments: expected Int32Array columns");
        return NULL;
    }
    napi_value a_sn = new_typed(env, napi_int32_array, 4, n, &sn);
    napi_value a_tm = new_typed(env, napi_float64_array, 8, n, &tm);
    napi_value a_st = new_typed(env, napi_uint8_array, 1, n, &st);
    if (!a_sn || !a_tm || !a_st) return NULL;
    if (hlsmap_views_from_fragments((hlsmap_map *)map, (int64_t)n, (const int32_t *)pl, (const int32_t *)lv,
                                    (const int32_t *)ix, (int32_t *)sn, (double *)tm, (uint8_t *)st) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_object(env, &out));
    napi_set_named_property(env, out, "sn", a_sn);
    napi_set_named_property(env, out, "time", a_tm);
    napi_set_named_property(env, out, "status", a_st);
    return out;
}

/* viewToJson(ctx, level, urlId, sn, time|null) -> {offsets: Int32Array(n+1), chars: Uint8Array}
 * (JSON.stringify(segmentView), test/segment-view.js:5-11, batched) */
static napi_value ViewToJson(napi_env env, napi_callback_info info)
{
    size_t argc = 5, n, m;
    napi_value argv[5], out;
    void *ctx, *lv, *ui, *sn, *tm, *off, *chars;
    int64_t total = 0;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    if (!typed(env, argv[1], napi_int32_array, &lv, &n) || !typed(env, argv[2], napi_int32_array, &ui, &m) ||
        m != n || !typed(env, argv[3], napi_int32_array, &sn, &m) || m != n ||
        !typed(env, argv[4], napi_float64_array, &tm, &m) || (tm && m != n)) {
        napi_throw_type_error(env, NULL, "viewToJson: expected Int32Array columns and a Float64Array or null");
        return NULL;
    }
    napi_value a_off = new_typed(env, napi_int32_array, 4, n + 1, &off);
    napi_value a_ch = new_typed(env, napi_uint8_array, 1, 104 * n + 1, &chars);
    if (!a_off || !a_ch) return NULL;
    if (hlsmap_view_to_json((hlsmap_ctx *)ctx, (int64_t)n, (const int32_t *)lv, (const int32_t *)ui,
                            (const int32_t *)sn, (const double *)tm, (int32_t *)off, (char *)chars,
                            (int64_t)(104 * n + 1), &total) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_object(env, &out));
    napi_set_named_property(env, out, "offsets", a_off);
    napi_set_named_property(env, out, "chars", a_ch);
    return out;
}

/* viewFromJson(ctx, chars: Uint8Array, offsets: Int32Array) -> {level, urlId, sn, time, flags, status}
 * (new SegmentView(JSON.parse(text)), batched) */
static napi_value ViewFromJson(napi_env env, napi_callback_info info)
{
    size_t argc = 3, n, nc;
    napi_value argv[3], out;
    void *ctx, *ch, *off, *lv, *ui, *sn, *tm, *fl, *st;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    if (!typed(env, argv[1], napi_uint8_array, &ch, &nc) || !typed(env, argv[2], napi_int32_array, &off, &n) || n < 1) {
        napi_throw_type_error(env, NULL, "viewFromJson: expected a Uint8Array and an Int32Array of offsets");
        return NULL;
    }
    n -= 1;
    napi_value a_lv = new_typed(env, napi_int32_array, 4, n, &lv), a_ui = new_typed(env, napi_int32_array, 4, n, &ui);
    napi_value a_sn = new_typed(env, napi_int32_array, 4, n, &sn), a_tm = new_typed(env, napi_float64_array, 8, n, &tm);
    napi_value a_fl = new_typed(env, napi_uint8_array, 1, n, &fl), a_st = new_typed(env, napi_uint8_array, 1, n, &st);
    if (!a_lv || !a_ui || !a_sn || !a_tm || !a_fl || !a_st) return NULL;
    if (hlsmap_view_from_json((hlsmap_ctx *)ctx, (int64_t)n, (const char *)ch, (const int32_t *)off, (int32_t *)lv,
                              (int32_t *)ui, (int32_t *)sn, (double *)tm, (uint8_t *)fl, (uint8_t *)st) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_object(env, &out));
    napi_set_named_property(env, out, "level", a_lv);
    napi_set_named_property(env, out, "urlId", a_ui);
    napi_set_named_property(env, out, "sn", a_sn);
    napi_set_named_property(env, out, "time", a_tm);
    napi_set_named_property(env, out, "flags", a_fl);
    napi_set_named_property(env, out, "status", a_st);
    return out;
}

static napi_value Init(napi_env env, napi_value exports)
{
    static const struct { const char *name; napi_callback fn; } table[] = {
        {"createContext", CreateContext}, {"createMap", CreateMap}, {"destroyMap", DestroyMap},
        {"getSegmentList", GetSegmentList}, {"getSegmentDuration", GetSegmentDuration},
        {"getTrackList", GetTrackList}, {"viewPack", ViewPack}, {"viewUnpack", ViewUnpack},
        {"viewIsEqual", ViewIsEqual}, {"viewFormat", ViewFormat}, {"createMapFromM3u8", CreateMapFromM3u8},
        {"viewsFromFragments", ViewsFromFragments}, {"viewToJson", ViewToJson}, {"viewFromJson", ViewFromJson},
    };
    for (size_t i = 0; i < sizeof table / sizeof table[0]; i++) {
        napi_value fn;
        if (napi_create_function(env, table[i].name, NAPI_AUTO_LENGTH, table[i].fn, NULL, &fn) != napi_ok ||
            napi_set_named_property(env, exports, table[i].name, fn) != napi_ok) {
            napi_throw_error(env, NULL, "hlsmap addon: could not register a function");
            return NULL;
        }
    }
    return exports;
}

NAPI_MODULE(NODE_GYP_MODULE_NAME, Init)

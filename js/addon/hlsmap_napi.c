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

/* A map handle owns its device tables: destroyMap() frees them at once, and a MediaMap that is
 * simply dropped gives them back when the garbage collector finalizes the handle. */
typedef struct { hlsmap_map *map; } map_box;
static void free_map(napi_env env, void *p, void *hint)
{
    (void)env; (void)hint;
    map_box *b = (map_box *)p;
    if (b->map) hlsmap_map_destroy(b->map);
    free(b);
}
static napi_value wrap_map(napi_env env, hlsmap_map *map)
{
    napi_value out;
    map_box *b = (map_box *)malloc(sizeof *b);
    if (!b) {
        hlsmap_map_destroy(map);
        napi_throw_error(env, NULL, "hlsmap addon: out of memory");
        return NULL;
    }
    b->map = map;
    if (napi_create_external(env, b, free_map, NULL, &out) != napi_ok) {
        free_map(env, b, NULL);
        napi_throw_error(env, NULL, "hlsmap addon: napi_create_external failed");
        return NULL;
    }
    return out;
}
/* -> the map behind a handle, or NULL (with a pending exception) for a destroyed / foreign one */
static hlsmap_map *get_map(napi_env env, napi_value v)
{
    void *p = NULL;
    if (napi_get_value_external(env, v, &p) != napi_ok || !p || !((map_box *)p)->map) {
        napi_throw_type_error(env, NULL, "hlsmap addon: not a live map handle");
        return NULL;
    }
    return ((map_box *)p)->map;
}

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
    (void)out;
    return wrap_map(env, map);
}

static napi_value DestroyMap(napi_env env, napi_callback_info info)
{
    size_t argc = 1;
    napi_value argv[1];
    void *p = NULL;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &p));
    if (p && ((map_box *)p)->map) {
        hlsmap_map_destroy(((map_box *)p)->map);
        ((map_box *)p)->map = NULL;          /* the finalizer later frees only the box */
    }
    return NULL;
}

/* getSegmentList(map, playlist|null, level, begin, duration)
 *   -> {offsets, sn, time, status, totalHits, nUnparsed, nMissing}   (media-map.js:27-54, batched) */
static napi_value GetSegmentList(napi_env env, napi_callback_info info)
{
    size_t argc = 5, nq, n;
    napi_value argv[5], out, v;
    hlsmap_map *map;
    void *pl, *lv, *bg, *du, *off, *st, *sn, *tm;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    if (!(map = get_map(env, argv[0]))) return NULL;
    size_t npl = 0;
    if (!typed(env, argv[1], napi_int32_array, &pl, &npl) || !typed(env, argv[2], napi_int32_array, &lv, &nq) ||
        (pl && npl != nq) ||
        !typed(env, argv[3], napi_float64_array, &bg, &n) || n != nq ||
        !typed(env, argv[4], napi_float64_array, &du, &n) || n != nq) {
        napi_throw_type_error(env, NULL, "getSegmentList: expected Int32Array level (and playlist, or null), "
                                         "Float64Array begin/duration, all of one length");
        return NULL;
    }
    napi_value a_off = new_typed(env, napi_int32_array, 4, nq + 1, &off);
    napi_value a_st = new_typed(env, napi_uint8_array, 1, nq, &st);
    if (!a_off || !a_st) return NULL;
    hlsmap_result res;
    /* two-call pattern: count, then fill buffers of exactly the needed size */
    if (hlsmap_get_segment_list(map, (int64_t)nq, (const int32_t *)pl, (const int32_t *)lv,
                                (const double *)bg, (const double *)du, (int32_t *)off, NULL, NULL, 0,
                                (uint8_t *)st, &res) != HLSMAP_OK)
        return fail_hlsmap(env);
    napi_value a_sn = new_typed(env, napi_int32_array, 4, (size_t)res.total_hits, &sn);
    napi_value a_tm = new_typed(env, napi_float64_array, 8, (size_t)res.total_hits, &tm);
    if (!a_sn || !a_tm) return NULL;
    if (res.total_hits > 0 &&
        hlsmap_get_segment_list(map, (int64_t)nq, (const int32_t *)pl, (const int32_t *)lv,
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
    hlsmap_map *map;
    void *pl, *lv, *o, *st;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    if (!(map = get_map(env, argv[0]))) return NULL;
    if (!typed(env, argv[1], napi_int32_array, &pl, &n) || !typed(env, argv[2], napi_int32_array, &lv, &nq) ||
        (pl && n != nq)) {
        napi_throw_type_error(env, NULL, "getSegmentDuration: expected Int32Array level (and playlist of the same length, or null)");
        return NULL;
    }
    napi_value a_o = new_typed(env, napi_float64_array, 8, nq, &o);
    napi_value a_st = new_typed(env, napi_uint8_array, 1, nq, &st);
    if (!a_o || !a_st) return NULL;
    if (hlsmap_get_segment_duration(map, (int64_t)nq, (const int32_t *)pl, (const int32_t *)lv,
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
    hlsmap_map *map;
    void *lv, *ui;
    int32_t playlist = 0;
    int64_t n = 0;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    if (!(map = get_map(env, argv[0]))) return NULL;
    if (argc > 1) napi_get_value_int32(env, argv[1], &playlist);
    if (hlsmap_get_track_list(map, playlist, NULL, NULL, 0, &n) != HLSMAP_OK) return fail_hlsmap(env);
    napi_value a_lv = new_typed(env, napi_int32_array, 4, (size_t)n, &lv);
    napi_value a_ui = new_typed(env, napi_int32_array, 4, (size_t)n, &ui);
    if (!a_lv || !a_ui) return NULL;
    if (n > 0 && hlsmap_get_track_list(map, playlist, (int32_t *)lv, (int32_t *)ui, n, &n) != HLSMAP_OK)
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
    (void)out;
    return wrap_map(env, map);
}

/* viewsFromFragments(map, playlist|null, level, index) -> {sn, time, status}
 * (`new SegmentView({sn: frag.sn, trackView, time: frag.start})`, p2p-loader-generator.js:145-147, batched) */
static napi_value ViewsFromFragments(napi_env env, napi_callback_info info)
{
    size_t argc = 4, n, m;
    napi_value argv[4], out;
    hlsmap_map *map;
    void *pl, *lv, *ix, *sn, *tm, *st;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    if (!(map = get_map(env, argv[0]))) return NULL;
    size_t npl = 0;
    if (!typed(env, argv[1], napi_int32_array, &pl, &npl) || !typed(env, argv[2], napi_int32_array, &lv, &n) ||
        (pl && npl != n) || !typed(env, argv[3], napi_int32_array, &ix, &m) || m != n) {
        napi_throw_type_error(env, NULL, "viewsFromFragments: expected Int32Array columns");
        return NULL;
    }
    napi_value a_sn = new_typed(env, napi_int32_array, 4, n, &sn);
    napi_value a_tm = new_typed(env, napi_float64_array, 8, n, &tm);
    napi_value a_st = new_typed(env, napi_uint8_array, 1, n, &st);
    if (!a_sn || !a_tm || !a_st) return NULL;
    if (hlsmap_views_from_fragments(map, (int64_t)n, (const int32_t *)pl, (const int32_t *)lv,
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

/* getSegmentRange(map, playlist|null, level, begin, duration) -> {first, count, status}
 * (media-map.js:27-54 where the caller only needs which fragments, not copies of them: 9 bytes
 * back per query instead of 12 per hit) */
static napi_value GetSegmentRange(napi_env env, napi_callback_info info)
{
    size_t argc = 5, nq, n, npl = 0;
    napi_value argv[5], out;
    hlsmap_map *map;
    void *pl, *lv, *bg, *du, *first, *count, *st;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    if (!(map = get_map(env, argv[0]))) return NULL;
    if (!typed(env, argv[1], napi_int32_array, &pl, &npl) || !typed(env, argv[2], napi_int32_array, &lv, &nq) ||
        (pl && npl != nq) || !typed(env, argv[3], napi_float64_array, &bg, &n) || n != nq ||
        !typed(env, argv[4], napi_float64_array, &du, &n) || n != nq) {
        napi_throw_type_error(env, NULL, "getSegmentRange: expected Int32Array level (and playlist, or null), "
                                         "Float64Array begin/duration, all of one length");
        return NULL;
    }
    napi_value a_f = new_typed(env, napi_int32_array, 4, nq, &first);
    napi_value a_c = new_typed(env, napi_int32_array, 4, nq, &count);
    napi_value a_s = new_typed(env, napi_uint8_array, 1, nq, &st);
    if (!a_f || !a_c || !a_s) return NULL;
    if (hlsmap_get_segment_range(map, (int64_t)nq, (const int32_t *)pl, (const int32_t *)lv, (const double *)bg,
                                 (const double *)du, (int32_t *)first, (int32_t *)count, (uint8_t *)st) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_object(env, &out));
    napi_set_named_property(env, out, "first", a_f);
    napi_set_named_property(env, out, "count", a_c);
    napi_set_named_property(env, out, "status", a_s);
    return out;
}

/* slideTracks(map, {playlist|null, level, drop, appendOffset: BigInt64Array(n+1), newStart, newSn, newDur0|null})
 * A playlist reload as the reference sees it (it reads hls.levels[..].details.fragments on every
 * call, media-map.js:28,41): per listed level, evict `drop` fragments at the front and append the new
 * ones, in place on the GPU.  Returns {allMonotone} so the caller knows which kernel now serves the map. */
static napi_value SlideTracks(napi_env env, napi_callback_info info)
{
    size_t argc = 2, n, m, nn, npl = 0;
    napi_value argv[2], out, v;
    hlsmap_map *map;
    void *pl, *lv, *dr, *ao, *ns, *nsn, *nd;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    if (!(map = get_map(env, argv[0]))) return NULL;
    if (!typed(env, field(env, argv[1], "playlist"), napi_int32_array, &pl, &npl) ||
        !typed(env, field(env, argv[1], "level"), napi_int32_array, &lv, &n) || (pl && npl != n) ||
        !typed(env, field(env, argv[1], "drop"), napi_int32_array, &dr, &m) || m != n ||
        !typed(env, field(env, argv[1], "appendOffset"), napi_bigint64_array, &ao, &m) || m != n + 1 ||
        !typed(env, field(env, argv[1], "newStart"), napi_float64_array, &ns, &nn) ||
        !typed(env, field(env, argv[1], "newSn"), napi_int32_array, &nsn, &m) || m != nn ||
        !typed(env, field(env, argv[1], "newDur0"), napi_float64_array, &nd, &m) || (nd && m != n) ||
        (n > 0 && (int64_t)nn != ((const int64_t *)ao)[n])) {
        napi_throw_type_error(env, NULL, "slideTracks: expected {playlist|null, level, drop, appendOffset, newStart, newSn, newDur0|null}");
        return NULL;
    }
    if (hlsmap_map_slide_tracks(map, (int64_t)n, (const int32_t *)pl, (const int32_t *)lv, (const int32_t *)dr,
                                (const int64_t *)ao, (const double *)ns, (const int32_t *)nsn, (const double *)nd) != HLSMAP_OK)
        return fail_hlsmap(env);
    hlsmap_map_info mi;
    if (hlsmap_map_get_info(map, &mi) != HLSMAP_OK) return fail_hlsmap(env);
    CHECK(env, napi_create_object(env, &out));
    napi_create_double(env, (double)mi.all_monotone, &v);
    napi_set_named_property(env, out, "allMonotone", v);
    napi_create_double(env, (double)mi.n_fragments, &v);
    napi_set_named_property(env, out, "fragments", v);
    return out;
}

/* viewIsInTrack(ctx, level, urlId, trackLevel, trackUrlId) -> Uint8Array   (segment-view.js:45-47, batched) */
static napi_value ViewIsInTrack(napi_env env, napi_callback_info info)
{
    size_t argc = 5, n = 0, m;
    napi_value argv[5];
    void *ctx, *p[4], *o;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    for (int i = 0; i < 4; i++) {
        if (!typed(env, argv[1 + i], napi_int32_array, &p[i], &m) || (i && m != n)) {
            napi_throw_type_error(env, NULL, "viewIsInTrack: expected four Int32Array of equal length");
            return NULL;
        }
        n = m;
    }
    napi_value a = new_typed(env, napi_uint8_array, 1, n, &o);
    if (!a) return NULL;
    if (hlsmap_view_is_in_track((hlsmap_ctx *)ctx, (int64_t)n, (const int32_t *)p[0], (const int32_t *)p[1],
                                (const int32_t *)p[2], (const int32_t *)p[3], (uint8_t *)o) != HLSMAP_OK)
        return fail_hlsmap(env);
    return a;
}

/* viewRoundtrip(ctx, level, urlId, sn) -> {wire: Uint32Array(3n), equal: Uint8Array}
 * (toArrayBuffer -> fromArrayBuffer -> isEqual, test/segment-view.js:21-27, one fused pass) */
static napi_value ViewRoundtrip(napi_env env, napi_callback_info info)
{
    size_t argc = 4, n, m;
    napi_value argv[4], out;
    void *ctx, *lv, *ui, *sn, *wire, *eq;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    if (!typed(env, argv[1], napi_int32_array, &lv, &n) || !typed(env, argv[2], napi_int32_array, &ui, &m) ||
        m != n || !typed(env, argv[3], napi_int32_array, &sn, &m) || m != n) {
        napi_throw_type_error(env, NULL, "viewRoundtrip: expected three Int32Array of equal length");
        return NULL;
    }
    napi_value a_w = new_typed(env, napi_uint32_array, 4, 3 * n, &wire);
    napi_value a_e = new_typed(env, napi_uint8_array, 1, n, &eq);
    if (!a_w || !a_e) return NULL;
    if (hlsmap_view_roundtrip((hlsmap_ctx *)ctx, (int64_t)n, (const int32_t *)lv, (const int32_t *)ui,
                              (const int32_t *)sn, (uint32_t *)wire, (uint8_t *)eq) != HLSMAP_OK)
        return fail_hlsmap(env);
    CHECK(env, napi_create_object(env, &out));
    napi_set_named_property(env, out, "wire", a_w);
    napi_set_named_property(env, out, "equal", a_e);
    return out;
}

/* hostAlloc(ctx, bytes) -> ArrayBuffer over pinned host memory (hlsmap_host_alloc): typed arrays
 * built on it travel to and from the GPU without a staging copy.  Freed (hlsmap_host_free) when the
 * ArrayBuffer is collected; the context must still be alive then. */
static void free_pinned(napi_env env, void *p, void *hint) { (void)env; hlsmap_host_free((hlsmap_ctx *)hint, p); }
static napi_value HostAlloc(napi_env env, napi_callback_info info)
{
    size_t argc = 2;
    napi_value argv[2], out;
    void *ctx, *p = NULL;
    int64_t bytes = 0;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    if (argc < 2 || napi_get_value_int64(env, argv[1], &bytes) != napi_ok || bytes < 0) {
        napi_throw_type_error(env, NULL, "hostAlloc: expected a byte count");
        return NULL;
    }
    if (hlsmap_host_alloc((hlsmap_ctx *)ctx, bytes, &p) != HLSMAP_OK) return fail_hlsmap(env);
    if (napi_create_external_arraybuffer(env, p, (size_t)bytes, free_pinned, ctx, &out) != napi_ok) {
        hlsmap_host_free((hlsmap_ctx *)ctx, p);
        napi_throw_error(env, NULL, "hlsmap addon: this runtime does not allow external ArrayBuffers");
        return NULL;
    }
    return out;
}

/* setBatchOverlap(ctx, enable): hlsmap_ctx_set_batch_overlap */
static napi_value SetBatchOverlap(napi_env env, napi_callback_info info)
{
    size_t argc = 2;
    napi_value argv[2];
    void *ctx;
    bool on = false;
    CHECK(env, napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
    CHECK(env, napi_get_value_external(env, argv[0], &ctx));
    if (argc > 1) napi_get_value_bool(env, argv[1], &on);
    if (hlsmap_ctx_set_batch_overlap((hlsmap_ctx *)ctx, on ? 1 : 0) != HLSMAP_OK) return fail_hlsmap(env);
    return NULL;
}

static napi_value Init(napi_env env, napi_value exports)
{
    static const struct { const char *name; napi_callback fn; } table[] = {
        {"createContext", CreateContext}, {"createMap", CreateMap}, {"destroyMap", DestroyMap},
        {"getSegmentList", GetSegmentList}, {"getSegmentDuration", GetSegmentDuration},
        {"getTrackList", GetTrackList}, {"viewPack", ViewPack}, {"viewUnpack", ViewUnpack},
        {"viewIsEqual", ViewIsEqual}, {"viewFormat", ViewFormat}, {"createMapFromM3u8", CreateMapFromM3u8},
        {"viewsFromFragments", ViewsFromFragments}, {"viewToJson", ViewToJson}, {"viewFromJson", ViewFromJson},
        {"getSegmentRange", GetSegmentRange}, {"slideTracks", SlideTracks}, {"viewIsInTrack", ViewIsInTrack},
        {"viewRoundtrip", ViewRoundtrip}, {"hostAlloc", HostAlloc}, {"setBatchOverlap", SetBatchOverlap},
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

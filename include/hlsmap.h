/*
 * hlsmap.h -- C ABI of the B200-native MediaMap batch engine (libhlsmap.so).
 *
 * This is the drop-in boundary for ONE path of streamroot-hlsjs-p2p-wrapper: the
 * MediaMap / SegmentView / TrackView mapping layer (lib/integration/mapping/, three .js).  The
 * reference has no FFI of its own (it is pure JavaScript); the entry points below are
 * what an N-API addon bound under the reference's JS classes calls -- see
 * INTEGRATION.md for that stub.  Every function cites the reference code it replaces
 * (paths relative to the reference checkout).
 *
 * Conventions
 *   - plain C types only; no exceptions cross the boundary;
 *   - every function returns an int status: 0 = OK, negative = misuse / CUDA failure, and
 *     hlsmap_last_error() then returns a thread-local message;
 *   - per-query outcomes (level missing / not parsed yet) are DATA (HLSMAP_ST_*), not
 *     errors: the JS binding turns them into the reference's throw / console.warn;
 *   - `*_dev` variants take DEVICE pointers and a cudaStream_t (as void*) and do no
 *     host<->device copies; a NULL stream means the context's own stream (pass
 *     cudaStreamLegacy / cudaStreamPerThread to name CUDA's default streams); the
 *     un-suffixed variants take HOST pointers, copy in and out, and block until the result
 *     is in the caller's buffers;
 *   - a context is bound to one GPU and used by one thread at a time; batches issued
 *     through one context share its scratch memory and must not overlap on the GPU;
 *   - there is no CPU fallback: without a usable CUDA device every call fails.
 */
#ifndef HLSMAP_H
#define HLSMAP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HLSMAP_OK 0
#define HLSMAP_ERR_INVALID (-1)   /* bad argument */
#define HLSMAP_ERR_CUDA (-2)      /* CUDA runtime failure (message has the detail) */
#define HLSMAP_ERR_NOMEM (-3)
#define HLSMAP_ERR_CAPACITY (-4)  /* hit buffers too small; result.total_hits = needed */
#define HLSMAP_ERR_OVERFLOW (-5)  /* more than INT32_MAX hits in one batch: split it */
#define HLSMAP_ERR_UNSUPPORTED (-6) /* parse_m3u8: a number the parser cannot convert exactly */

/* per-query status */
#define HLSMAP_ST_OK 0
#define HLSMAP_ST_UNPARSED 1 /* level has no details: warn + [] (media-map.js:34-37) */
#define HLSMAP_ST_MISSING 2  /* no such level: the reference throws (media-map.js:30-32) */
#define HLSMAP_ST_EMPTY 3    /* getSegmentDuration on an empty list (media-map.js:86) */
#define HLSMAP_ST_SCATTERED 4 /* range form only: hits are not one contiguous run */
#define HLSMAP_ST_NO_FRAGMENT 5 /* views_from_fragments: fragments[index] is undefined */

/* track_state values for hlsmap_map_create */
#define HLSMAP_TRACK_MISSING 0  /* hls.levels[i] is falsy */
#define HLSMAP_TRACK_UNPARSED 1 /* level.details is falsy */
#define HLSMAP_TRACK_PARSED 2

/* resolve kernel selection (hlsmap_map_set_path) */
#define HLSMAP_PATH_AUTO 0
#define HLSMAP_PATH_SEARCH 1 /* binary search; only legal when every track is monotone */
#define HLSMAP_PATH_FILTER 2 /* the reference's linear filter, any table */

typedef struct hlsmap_ctx hlsmap_ctx; /* one GPU: streams, scratch, staging buffers */
typedef struct hlsmap_map hlsmap_map; /* uploaded fragment tables of many playlists */

typedef struct hlsmap_result {
    int64_t total_hits;      /* H: SegmentViews the batch produced */
    int64_t n_unparsed;      /* queries with HLSMAP_ST_UNPARSED */
    int64_t n_missing;       /* queries with HLSMAP_ST_MISSING  */
    int32_t kernel_launches; /* CUDA kernels this call launched */
    int32_t path;            /* HLSMAP_PATH_SEARCH or HLSMAP_PATH_FILTER actually run */
} hlsmap_result;

typedef struct hlsmap_map_info {
    int32_t n_playlists;
    int32_t n_tracks;
    int64_t n_fragments;
    int32_t n_monotone_tracks; /* parsed tracks whose start column is sorted and NaN-free */
    int32_t n_parsed_tracks;
    int32_t all_monotone;      /* 1 => the search kernel is exact for every track */
    int32_t stage_all;         /* 1 => the whole start column is staged in shared memory */
    int64_t device_bytes;      /* HBM held by this map */
} hlsmap_map_info;

/* ---- library / context ------------------------------------------------------------ */
const char *hlsmap_last_error(void);
int hlsmap_version(void);
/* Number of CUDA devices visible, or a negative error. */
int hlsmap_device_count(void);

int hlsmap_ctx_create(int device, hlsmap_ctx **out);
void hlsmap_ctx_destroy(hlsmap_ctx *ctx);
int hlsmap_ctx_device(const hlsmap_ctx *ctx);
/* Opt in to overlapping consecutive getSegmentList batches issued with the _dev entry point:
 * with enable != 0 a batch may begin (programmatic dependent launch) while the previous batch
 * on the same stream is still writing its last hits, which hides each batch's ramp-up behind
 * its predecessor's tail.  ONLY legal when consecutive batches are independent: the next batch
 * must not read, or write to, any buffer the previous one is still writing.  Off by default
 * (plain stream order).  The library's own scratch is double-buffered for this.  With enable = 1
 * hlsmap_get_segment_list_dev_batches additionally resolves runs of batches with one launch (see
 * there); enable = 2 keeps one (overlapping) launch per batch. */
int hlsmap_ctx_set_batch_overlap(hlsmap_ctx *ctx, int enable);
/* Block until everything issued through the context's own stream has finished. */
int hlsmap_ctx_synchronize(hlsmap_ctx *ctx);

/* Pinned host memory for bindings that want zero-staging copies (N-API external
 * ArrayBuffers, the Python test binding).  Optional: every host-pointer entry point also
 * accepts pageable memory. */
int hlsmap_host_alloc(hlsmap_ctx *ctx, int64_t bytes, void **out);
int hlsmap_host_free(hlsmap_ctx *ctx, void *p);

/* ---- table upload: `new MediaMap(hls)` (media-map.js:6-8) ------------------------- */
/* The reference keeps a live reference to hls.levels; here the level tables of
 * n_playlists players are copied to HBM as SoA columns (input shape: media-map.js:28,41,
 * 43-48,61,66-69,82-84).
 *   playlist_track_base[n_playlists+1]  first track of each playlist; a playlist's
 *                                       `levels.length` is base[p+1]-base[p]
 *   track_offset[n_tracks+1]            first fragment of each track (monotone, [0]=0)
 *   track_state[n_tracks]               HLSMAP_TRACK_*
 *   url_count[n_tracks]                 level.url.length (getTrackList), may be NULL (=1)
 *   start[F], sn[F]                     fragment.start / fragment.sn in index order
 *   duration[F]                         fragment.duration, may be NULL (then NaN)
 * Monotonicity of every start column is established on the GPU at upload. */
int hlsmap_map_create(hlsmap_ctx *ctx, int32_t n_playlists, const int32_t *playlist_track_base,
                      int32_t n_tracks, const int64_t *track_offset, const uint8_t *track_state,
                      const int32_t *url_count, const double *start, const int32_t *sn,
                      const double *duration, hlsmap_map **out);
void hlsmap_map_destroy(hlsmap_map *map);
int hlsmap_map_get_info(const hlsmap_map *map, hlsmap_map_info *info);
/* Per-track monotone flags (1/0), n_tracks bytes; what routed AUTO to search or filter. */
int hlsmap_map_get_monotone(const hlsmap_map *map, uint8_t *flags);
/* Force a kernel (tests, benchmarks).  SEARCH on a map that is not all-monotone fails. */
int hlsmap_map_set_path(hlsmap_map *map, int path);

/* ---- live playlists: slide fragment windows in place ------------------------------- */
/* The reference holds a live reference to hls.levels (media-map.js:6-8) and reads
 * level.details.fragments on every call (:28,41), so a playlist reload is seen at once.  The
 * batch analogue: for each i, track (playlist[i], level[i]) evicts its first drop[i] fragments
 * and appends new_start/new_sn[append_offset[i] .. append_offset[i+1]) (append_offset[0] = 0);
 * new_dur0[i] becomes the duration of the track's new first fragment (getSegmentDuration);
 * the array may be NULL only when no listed track changes its first fragment (every drop[i] is 0
 * and nothing is appended to an empty track), otherwise HLSMAP_ERR_INVALID.  A track may appear once per call and may not outgrow the fragment
 * count it had at hlsmap_map_create (HLSMAP_ERR_CAPACITY).  Monotonicity is re-proven on the
 * GPU for every touched track; a track that stops being sorted routes the map to the filter
 * kernel.  Host pointers; blocks until the tables are updated. */
int hlsmap_map_slide_tracks(hlsmap_map *map, int64_t n, const int32_t *playlist, const int32_t *level,
                            const int32_t *drop, const int64_t *append_offset, const double *new_start,
                            const int32_t *new_sn, const double *new_dur0);

/* ---- getSegmentList (media-map.js:27-54), batched ---------------------------------- */
/* Query q = (playlist[q], level[q], begin[q], duration[q]); playlist may be NULL (all 0).
 * `urlId` never takes part in the lookup (media-map.js:28) -- bindings echo it.
 * List form: offsets[nq+1] is the exclusive prefix of hit counts (offsets[nq] = H); the
 * hits of query q are sn_out/time_out[offsets[q] .. offsets[q+1]) in fragment index
 * order: sn = fragment.sn (media-map.js:46), time = fragment.start bit-for-bit (:48).
 * Keeps fragment i iff begin <= start[i] && start[i] <= fl(begin + duration) (:44).
 * status (nullable) receives HLSMAP_ST_* per query.  If H > hit_capacity nothing beyond
 * the capacity is written and HLSMAP_ERR_CAPACITY is returned with result->total_hits
 * set (call again with larger buffers; sn_out/time_out may be NULL with capacity 0 to
 * size them).  result may be NULL for the _dev variant: the call is then fully
 * asynchronous and hlsmap_ctx_last_result() retrieves the counters later. */
int hlsmap_get_segment_list_dev(hlsmap_map *map, int64_t nq, const int32_t *d_playlist,
                                const int32_t *d_level, const double *d_begin,
                                const double *d_duration, int32_t *d_offsets, int32_t *d_sn_out,
                                double *d_time_out, int64_t hit_capacity, uint8_t *d_status,
                                void *stream, hlsmap_result *result);
int hlsmap_get_segment_list(hlsmap_map *map, int64_t nq, const int32_t *playlist,
                            const int32_t *level, const double *begin, const double *duration,
                            int32_t *offsets, int32_t *sn_out, double *time_out,
                            int64_t hit_capacity, uint8_t *status, hlsmap_result *result);
/* The same batch for callers that may not allocate.  (No counterpart in the reference: a
 * JavaScript library never sees an allocator; this exists for native hosts of the engine.)
 * Some callers cannot call cudaMalloc while they run: a real-time thread with a no-malloc
 * rule, a plugin inside a host process that owns all device memory and forbids allocation
 * from callbacks, a service whose own pool has already taken every byte of HBM.  The
 * regular _dev entry point grows its look-back scratch with cudaMalloc when a batch is
 * larger than any before it; this one never allocates or frees device memory, and on the
 * success path no host memory either (only a failing call builds its error message in a host
 * string).  Inputs and outputs are still the caller's device buffers; the scratch is a
 * fixed array inside the library image (about 0.5 MiB held by every GPU that loads the
 * library, until exit).
 *   - Limit: 65536 tiles per call.  A tile holds 256 to 4096 queries depending on the map
 *     and the kernel that serves it (at least 16 777 216 queries per call, 33 554 432 or more
 *     on sorted maps).  Larger batches are refused with HLSMAP_ERR_INVALID and a message that
 *     says how many queries fit on this map -- split the batch.
 *   - The first call on each GPU looks up the array's device address; make that first
 *     call before starting any stream capture.
 *   - All calls share the one array: calls that overlap on the GPU (two streams, two
 *     threads) corrupt each other's offsets.  Issue them in stream order or serialise.
 *   - The array keeps the previous call's contents; every call zeroes the words it reads.
 * result == NULL keeps the call asynchronous (the counters are then not retrievable: use
 * the offsets[nq] entry for the hit total). */
int hlsmap_get_segment_list_noalloc_dev(hlsmap_map *map, int64_t nq, const int32_t *d_playlist,
                                        const int32_t *d_level, const double *d_begin,
                                        const double *d_duration, int32_t *d_offsets,
                                        int32_t *d_sn_out, double *d_time_out,
                                        int64_t hit_capacity, uint8_t *d_status, void *stream,
                                        hlsmap_result *result);
/* Several independent list-form batches submitted with ONE call (what a serving loop does with the
 * requests that have queued up): the same results as calling hlsmap_get_segment_list_dev for
 * batches[0..n) in order with result = NULL, without returning to the host language in between.
 * With hlsmap_ctx_set_batch_overlap (the batches must not share output buffers) runs of up to 32
 * non-empty batches are resolved by ONE kernel launch each on maps that take the search path --
 * their tiles are handed out in batch order, every batch keeps its own offsets, capacity and status
 * column, and the tables are staged once per CTA for the whole run -- so n batches pay one ramp-up
 * and one tail instead of n (10^6-query batches on a B200: 30.6 us per batch against 33.2 as
 * overlapping launches); without that option, one launch per batch.  Stops at the first batch
 * that cannot be launched and returns its error; *n_launched (nullable) = batches issued (a run
 * ends in front of such a batch, so everything before it has been issued).  Afterwards
 * hlsmap_ctx_last_result reports the last batch's counters and, in kernel_launches, the launches
 * of the whole call.  The reference has one call per query (media-map.js:27); this is that call
 * over many query batches. */
typedef struct hlsmap_batch {
    int64_t nq;
    const int32_t *d_playlist;   /* nullable */
    const int32_t *d_level;
    const double *d_begin;
    const double *d_duration;
    int32_t *d_offsets;          /* [nq + 1] */
    int32_t *d_sn_out;           /* nullable together with d_time_out (count only) */
    double *d_time_out;
    int64_t hit_capacity;
    uint8_t *d_status;           /* nullable */
} hlsmap_batch;
int hlsmap_get_segment_list_dev_batches(hlsmap_map *map, int32_t n, const hlsmap_batch *batches,
                                        void *stream, int32_t *n_launched);
/* The same, bracketed by the caller's CUDA events: event_begin (a cudaEvent_t, nullable) is recorded
 * on the stream immediately before the first launch and event_end (nullable) immediately after the
 * last, from inside the call -- so that the device time of the n batches can be measured without
 * the host language's call overhead between "event recorded" and "first kernel issued". */
int hlsmap_get_segment_list_dev_batches_timed(hlsmap_map *map, int32_t n, const hlsmap_batch *batches,
                                              void *stream, void *event_begin, void *event_end,
                                              int32_t *n_launched);
/* Counters of the most recent batch issued through ctx (synchronises its stream). */
int hlsmap_ctx_last_result(hlsmap_ctx *ctx, void *stream, hlsmap_result *result);

/* Range form: first[q] = index (within its track) of the first kept fragment or -1,
 * count[q] = number kept.  Exact only where the kept fragments are contiguous: a
 * non-monotone track yields HLSMAP_ST_SCATTERED with the true count and first hit. */
int hlsmap_get_segment_range_dev(hlsmap_map *map, int64_t nq, const int32_t *d_playlist,
                                 const int32_t *d_level, const double *d_begin,
                                 const double *d_duration, int32_t *d_first, int32_t *d_count,
                                 uint8_t *d_status, void *stream);
int hlsmap_get_segment_range(hlsmap_map *map, int64_t nq, const int32_t *playlist,
                             const int32_t *level, const double *begin, const double *duration,
                             int32_t *first, int32_t *count, uint8_t *status);

/* ---- getSegmentDuration (media-map.js:81-87), batched ------------------------------ */
/* out[q] = duration of the FIRST fragment of the level, whatever the view's sn. */
int hlsmap_get_segment_duration_dev(hlsmap_map *map, int64_t n, const int32_t *d_playlist,
                                    const int32_t *d_level, double *d_out, uint8_t *d_status,
                                    void *stream);
int hlsmap_get_segment_duration(hlsmap_map *map, int64_t n, const int32_t *playlist,
                                const int32_t *level, double *out, uint8_t *status);

/* ---- SegmentView construction from fragments (the loader / track-change call sites) --- */
/* `new SegmentView({sn: frag.sn, trackView: {level, urlId: level.urlId}, time: frag.start})`
 * (p2p-loader-generator.js:145-147; TrackView on a level switch: player-interface.js:15-20).
 * For fragment index[q] of (playlist[q], level[q]): sn_out[q] = fragment.sn, time_out[q] =
 * fragment.start.  level and urlId are the caller's own columns (urlId is hls.js state the
 * tables do not hold), so they are not copied.  status: MISSING / UNPARSED / NO_FRAGMENT. */
int hlsmap_views_from_fragments_dev(hlsmap_map *map, int64_t n, const int32_t *d_playlist,
                                    const int32_t *d_level, const int32_t *d_index, int32_t *d_sn_out,
                                    double *d_time_out, uint8_t *d_status, void *stream);
int hlsmap_views_from_fragments(hlsmap_map *map, int64_t n, const int32_t *playlist,
                                const int32_t *level, const int32_t *index, int32_t *sn_out,
                                double *time_out, uint8_t *status);

/* ---- getNextSegmentView: EXTENSION -------------------------------------------------- */
/* BASELINE.json configs[2] names getNextSegmentView, but this reference does not have it: it
 * was removed in 3.2.5 (CHANGELOG.md:89) and no implementation or test survives, so there is
 * nothing to be bit-exact with.  This entry point is the library's own, documented definition:
 * in level (playlist[q], level[q]), take the first fragment whose sn equals sn[q]; if a
 * fragment follows it, found[q] = 1 and sn_out/time_out[q] are that next fragment's sn and
 * start; otherwise found[q] = 0.  Parity is claimed against oracle/ only. */
int hlsmap_get_next_segment_view_dev(hlsmap_map *map, int64_t n, const int32_t *d_playlist,
                                     const int32_t *d_level, const int32_t *d_sn, int32_t *d_sn_out,
                                     double *d_time_out, uint8_t *d_found, uint8_t *d_status,
                                     void *stream);
int hlsmap_get_next_segment_view(hlsmap_map *map, int64_t n, const int32_t *playlist,
                                 const int32_t *level, const int32_t *sn, int32_t *sn_out,
                                 double *time_out, uint8_t *found, uint8_t *status);

/* ---- JSON form of a SegmentView (test/segment-view.js:5-11) ------------------ */
/* JSON.stringify(segmentView): the object the constructor built (segment-view.js:22-26 -- sn, then
 * trackView = TrackView{level, urlId} (track-view.js:6-9), then time) walked in property order,
 *     {"sn":25,"trackView":{"level":1,"urlId":0},"time":12.5}
 * numbers printed by ECMAScript Number::toString (shortest digits that read back as the same
 * double, four layouts by exponent), a NaN / +-Infinity time printed as null, and -- for a view
 * with no time (time == NULL here), e.g. one made by fromArrayBuffer -- the property left out.
 * offsets [n + 1] out; chars [capacity] out; *total = characters needed (HLSMAP_ERR_CAPACITY when
 * capacity is smaller; offsets are valid then).  At most 104 characters per view. */
int hlsmap_view_to_json_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *d_level, const int32_t *d_url_id,
                            const int32_t *d_sn, const double *d_time, int32_t *d_offsets, char *d_chars,
                            int64_t capacity, void *stream, int64_t *total);
int hlsmap_view_to_json(hlsmap_ctx *ctx, int64_t n, const int32_t *level, const int32_t *url_id,
                        const int32_t *sn, const double *time, int32_t *offsets, char *chars,
                        int64_t capacity, int64_t *total);

/* new SegmentView(JSON.parse(text)) for texts of the shape JSON.stringify gives a SegmentView (keys
 * in constructor order, each optional, no white space).  flags[i] says which properties were there;
 * a missing one leaves 0 / NaN in its column, as `undefined` has no column value.  status[i]:
 * OK, MALFORMED (not that shape), UNSUPPORTED (a number the exact converter does not cover: decimal
 * exponent beyond +19 / -38 after 19 digits), NOT_INT32 (sn / level / urlId not an int32). */
#define HLSMAP_JSON_OK 0
#define HLSMAP_JSON_MALFORMED 1
#define HLSMAP_JSON_UNSUPPORTED 2
#define HLSMAP_JSON_NOT_INT32 3
#define HLSMAP_JSON_HAS_SN 1
#define HLSMAP_JSON_HAS_TRACKVIEW 2
#define HLSMAP_JSON_HAS_LEVEL 4
#define HLSMAP_JSON_HAS_URLID 8
#define HLSMAP_JSON_HAS_TIME 16
#define HLSMAP_JSON_TIME_NULL 32
int hlsmap_view_from_json_dev(hlsmap_ctx *ctx, int64_t n, const char *d_chars, const int32_t *d_offsets,
                              int32_t *d_level, int32_t *d_url_id, int32_t *d_sn, double *d_time,
                              uint8_t *d_flags, uint8_t *d_status, void *stream);
int hlsmap_view_from_json(hlsmap_ctx *ctx, int64_t n, const char *chars, const int32_t *offsets,
                          int32_t *level, int32_t *url_id, int32_t *sn, double *time, uint8_t *flags,
                          uint8_t *status);

/* ---- media playlists (M3U8 text) -> fragment tables --------------------- */
/* The step BEFORE the mapping layer.  The reference never parses a playlist itself: hls.js does
 * (package.json:34, "hls.js": "0.5.46", absent from the reference checkout), and media-map.js:28,
 * 41-48 reads the `level.details.fragments` it leaves behind.  These entry points replace that
 * hand-over for the batch case: the texts of many level playlists go to the GPU as bytes and come
 * back as the SoA fragment tables hlsmap_map_create() takes -- or go straight into a map.
 *
 * What is restated of hls.js 0.5.46's parseLevelPlaylist (src/loader/playlist-loader.js), the
 * part the mapping layer can observe: one global regular expression walks the text; of its
 * alternatives `#EXT-X-MEDIA-SEQUENCE:(\d+)` sets the next sn (parseInt), `#EXTINF:([\d.]+)...`
 * appends {sn: next sn++, start: running sum, duration: parseFloat(run)} unless that is NaN,
 * `#EXT-X-ENDLIST` clears `live`, and the alternatives that swallow the rest of their line
 * (EXTINF, `#EXT-X-KEY:(.*)`, `#EXT-X-PROGRAM-DATE-TIME:(.*)`) hide tags later on that line.
 * `start` is summed in playlist order in fp64, so it carries the reference's bits.
 * Parity: pinned only to the CPU restatement in oracle/ (glibc strtod as the independent decimal
 * conversion) -- the parser's own source and tests are not in the reference checkout.
 *
 * text         all playlists' bytes, concatenated; track t owns [text_offset[t], text_offset[t+1])
 * track_state  [n_tracks] HLSMAP_TRACK_*, or NULL (all parsed); the text of a track that is not
 *              HLSMAP_TRACK_PARSED is ignored and the track gets no fragments
 * track_offset [n_tracks + 1] out: first fragment of each track
 * track_live   [n_tracks] out: 1 unless the playlist holds #EXT-X-ENDLIST
 * start / sn / duration  [capacity] out, or all NULL with capacity 0 to only count
 * Numbers the converter cannot round exactly (see csrc/decimal.cuh: more than 19 significant digits
 * whose tail could change the rounding, or a sequence number beyond int32) are never guessed: the
 * call fails with HLSMAP_ERR_UNSUPPORTED and res->n_unsupported says how many there were.  The same
 * goes for a '#' more than 16384 bytes into its line: whether an earlier tag on the line hides it is
 * decided by replaying the line, and that work is bounded.
 * HLSMAP_ERR_CAPACITY: res->n_fragments entries are needed; track_offset / track_live are valid. */
typedef struct hlsmap_ingest_result {
    int64_t n_fragments;
    int64_t n_sequence_tags;
    int64_t n_unsupported;
    int64_t text_bytes;
    int32_t kernel_launches;
    int32_t reserved;
} hlsmap_ingest_result;

int hlsmap_parse_m3u8_dev(hlsmap_ctx *ctx, int32_t n_tracks, const uint8_t *d_track_state, const char *d_text,
                          const int64_t *d_text_offset, int64_t text_bytes, int64_t *d_track_offset,
                          uint8_t *d_track_live, double *d_start, int32_t *d_sn, double *d_duration,
                          int64_t capacity, hlsmap_ingest_result *res, void *stream);
int hlsmap_parse_m3u8(hlsmap_ctx *ctx, int32_t n_tracks, const uint8_t *track_state, const char *text,
                      const int64_t *text_offset, int64_t *track_offset, uint8_t *track_live, double *start,
                      int32_t *sn, double *duration, int64_t capacity, hlsmap_ingest_result *res);
/* new MediaMap(hls) for playlists still in text form: parse on the device, keep the columns there. */
int hlsmap_map_create_from_m3u8(hlsmap_ctx *ctx, int32_t n_playlists, const int32_t *playlist_track_base,
                                int32_t n_tracks, const uint8_t *track_state, const int32_t *url_count,
                                const char *text, const int64_t *text_offset, hlsmap_ingest_result *res,
                                hlsmap_map **out);
/* The same with the text and its offsets already in device memory (text_offset must start at 0, be
 * monotone and end at text_bytes: it is not checked on the host); the small per-track metadata stays
 * on the host.  Work is issued on the context's stream; the call returns when the map is ready. */
int hlsmap_map_create_from_m3u8_dev(hlsmap_ctx *ctx, int32_t n_playlists, const int32_t *playlist_track_base,
                                    int32_t n_tracks, const uint8_t *track_state, const int32_t *url_count,
                                    const char *d_text, const int64_t *d_text_offset, int64_t text_bytes,
                                    hlsmap_ingest_result *res, hlsmap_map **out);

/* ---- getTrackList (media-map.js:60-73) --------------------------------------------- */
/* Level-major, urlId-minor enumeration of one playlist's tracks.  Pass NULL outputs to
 * get the count in *n_out.  Host-side: it enumerates upload metadata, no arithmetic. */
int hlsmap_get_track_list(const hlsmap_map *map, int32_t playlist, int32_t *level_out,
                          int32_t *url_id_out, int64_t capacity, int64_t *n_out);

/* ---- SegmentView / TrackView value types over SoA columns -------------------------- */
/* A view is (level, urlId, sn, time) (segment-view.js:22-26, track-view.js:3-6). */

/* toArrayBuffer (segment-view.js:59-61): wire[3n] u32 = level, urlId, sn per view. */
int hlsmap_view_pack_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *d_level,
                         const int32_t *d_url_id, const int32_t *d_sn, uint32_t *d_wire,
                         void *stream);
int hlsmap_view_pack(hlsmap_ctx *ctx, int64_t n, const int32_t *level, const int32_t *url_id,
                     const int32_t *sn, uint32_t *wire);
/* fromArrayBuffer (segment-view.js:9-17): the parsed fields are Numbers in [0, 2^32). */
int hlsmap_view_unpack_dev(hlsmap_ctx *ctx, int64_t n, const uint32_t *d_wire, uint32_t *d_level,
                           uint32_t *d_url_id, uint32_t *d_sn, void *stream);
int hlsmap_view_unpack(hlsmap_ctx *ctx, int64_t n, const uint32_t *wire, uint32_t *level,
                       uint32_t *url_id, uint32_t *sn);
/* isEqual (segment-view.js:33-39, track-view.js:19-24): sn, level, urlId; not time. */
int hlsmap_view_is_equal_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *d_a_level,
                             const int32_t *d_a_url_id, const int32_t *d_a_sn,
                             const int32_t *d_b_level, const int32_t *d_b_url_id,
                             const int32_t *d_b_sn, uint8_t *d_out, void *stream);
int hlsmap_view_is_equal(hlsmap_ctx *ctx, int64_t n, const int32_t *a_level,
                         const int32_t *a_url_id, const int32_t *a_sn, const int32_t *b_level,
                         const int32_t *b_url_id, const int32_t *b_sn, uint8_t *out);
/* isInTrack (segment-view.js:45-47) == TrackView.isEqual (track-view.js:19-24). */
int hlsmap_view_is_in_track_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *d_level,
                                const int32_t *d_url_id, const int32_t *d_t_level,
                                const int32_t *d_t_url_id, uint8_t *d_out, void *stream);
int hlsmap_view_is_in_track(hlsmap_ctx *ctx, int64_t n, const int32_t *level,
                            const int32_t *url_id, const int32_t *t_level,
                            const int32_t *t_url_id, uint8_t *out);
/* toArrayBuffer -> fromArrayBuffer -> isEqual(original) fused (test/segment-view.js:21-27):
 * writes the 12-byte wire form and the equality flag (parse is done from registers). */
int hlsmap_view_roundtrip_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *d_level,
                              const int32_t *d_url_id, const int32_t *d_sn, uint32_t *d_wire,
                              uint8_t *d_equal, void *stream);
int hlsmap_view_roundtrip(hlsmap_ctx *ctx, int64_t n, const int32_t *level,
                          const int32_t *url_id, const int32_t *sn, uint32_t *wire,
                          uint8_t *equal);
/* viewToString (segment-view.js:52-54, track-view.js:11-13): "L{level}U{urlId}S{sn}"
 * (with_sn = 0: the TrackView form "L{level}U{urlId}").  offsets[n+1] are byte offsets
 * into chars; HLSMAP_ERR_CAPACITY (with *total_out = bytes needed) if it does not fit. */
int hlsmap_view_format_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *d_level,
                           const int32_t *d_url_id, const int32_t *d_sn, int with_sn,
                           int32_t *d_offsets, char *d_chars, int64_t capacity, void *stream,
                           int64_t *total_out);
int hlsmap_view_format(hlsmap_ctx *ctx, int64_t n, const int32_t *level, const int32_t *url_id,
                       const int32_t *sn, int with_sn, int32_t *offsets, char *chars,
                       int64_t capacity, int64_t *total_out);

#ifdef __cplusplus
}
#endif
#endif /* HLSMAP_H */

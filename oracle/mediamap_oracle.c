/*
 * mediamap_oracle.c -- CPU restatement of the reference's MediaMap / SegmentView /
 * TrackView path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the checker, never the product: only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The shipped path
 * (libhlsmap.so) never links or calls anything here.
 *
 * Parity pin: the reference is JavaScript and no JS engine exists in the build
 * container, so this restatement is pinned against the reference's own known-answer
 * tests, transcribed as data in tests/golden/reference_vectors.json
 * (test/media-map.js:8-137, test/segment-view.js:5-73, test/track-view.js:4-41, fixture
 * test/mocks/hls.js:12-45).  Functions with no reference test (getSegmentDuration, the
 * exact viewToString text) are restated from the source only and say so below.
 *
 * All `file:line` citations are relative to the reference checkout.
 *
 * Table model (unpadded SoA, the batch form of `hls.levels[..].details.fragments`,
 * consumers media-map.js:28,41,43-48):
 *   playlist_base[P+1]  first track index of each playlist; levels = base[p+1]-base[p]
 *   track_off[T+1]      first fragment index of each track
 *   track_state[T]      0 = level slot is empty (falsy), 1 = level has no `details`,
 *                       2 = parsed
 *   start[F], sn[F]     fragment.start (fp64) and fragment.sn (int32), index order
 *   dur0[T]             duration of the track's first fragment (media-map.js:83-85)
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define ST_OK 0
#define ST_UNPARSED 1 /* media-map.js:34-37: warn + [] */
#define ST_MISSING 2  /* media-map.js:30-32: throws "getSegmentList: level doesn't exist" */

/* ---- getSegmentList, one query: the linear filter of media-map.js:42-51 ---------- */
/* Keeps fragment i iff  beginTime <= start[i] && start[i] <= beginTime + duration
 * (media-map.js:44), both ends inclusive, the sum rounded once to fp64 as JS does.
 * `volatile` pins the sum to a real fp64 add (no x87 excess precision, no contraction).
 * Output order is fragment index order.  sn_out/time_out may be NULL to only count. */
int64_t oracle_filter_one(const double *start, const int32_t *sn, int64_t n, double begin,
                          double duration, int32_t *sn_out, double *time_out)
{
    volatile double upper = begin + duration;
    double ub = upper;
    int64_t h = 0;
    for (int64_t i = 0; i < n; i++) {
        double s = start[i];
        if (begin <= s && s <= ub) {
            if (sn_out) sn_out[h] = sn[i];     /* media-map.js:46 */
            if (time_out) time_out[h] = s;     /* media-map.js:48 */
            h++;
        }
    }
    return h;
}

/* Level lookup of media-map.js:28-37.  Returns the status and, when ST_OK, the track. */
static int resolve_track(int32_t n_playlists, const int32_t *playlist_base,
                         const uint8_t *track_state, int32_t playlist, int32_t level,
                         int32_t *track)
{
    if (playlist < 0 || playlist >= n_playlists) return ST_MISSING;
    int32_t b = playlist_base[playlist], e = playlist_base[playlist + 1];
    if (level < 0 || level >= e - b) return ST_MISSING; /* levels[level] is undefined */
    int32_t t = b + level;
    if (track_state[t] == 0) return ST_MISSING;          /* falsy slot, :30 */
    if (track_state[t] == 1) return ST_UNPARSED;         /* no details, :34 */
    *track = t;
    return ST_OK;
}

typedef struct {
    int32_t n_playlists;
    const int32_t *playlist_base;
    const int64_t *track_off;
    const uint8_t *track_state;
    const double *start;
    const int32_t *sn;
    int64_t q0, q1;
    const int32_t *playlist, *level;
    const double *begin, *duration;
    int64_t *offsets; /* pass 1: counts at [q+1]; pass 2: exclusive offsets */
    uint8_t *status;
    int32_t *sn_out;
    double *time_out;
    int64_t cap;
    int pass;
    uint64_t checksum;
    int64_t hits;
} job_t;

static void *batch_worker(void *arg)
{
    job_t *j = (job_t *)arg;
    for (int64_t q = j->q0; q < j->q1; q++) {
        int32_t t = 0;
        int st = resolve_track(j->n_playlists, j->playlist_base, j->track_state,
                               j->playlist ? j->playlist[q] : 0, j->level[q], &t);
        if (j->pass == 1) {
            if (j->status) j->status[q] = (uint8_t)st;
            int64_t c = 0;
            if (st == ST_OK) {
                int64_t o = j->track_off[t], n = j->track_off[t + 1] - o;
                c = oracle_filter_one(j->start + o, j->sn + o, n, j->begin[q], j->duration[q],
                                      NULL, NULL);
            }
            j->offsets[q + 1] = c;
        } else if (st == ST_OK) {
            int64_t o = j->track_off[t], n = j->track_off[t + 1] - o;
            int64_t at = j->offsets[q], c = j->offsets[q + 1] - at;
            if (c > 0 && at + c <= j->cap)
                oracle_filter_one(j->start + o, j->sn + o, n, j->begin[q], j->duration[q],
                                  j->sn_out + at, j->time_out + at);
        }
    }
    return NULL;
}

static void run_threads(job_t *proto, int64_t n_queries, int threads)
{
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t tid[256];
    job_t jobs[256];
    for (int i = 0; i < threads; i++) {
        jobs[i] = *proto;
        jobs[i].q0 = n_queries * i / threads;
        jobs[i].q1 = n_queries * (i + 1) / threads;
        if (threads == 1) { batch_worker(&jobs[i]); return; }
        pthread_create(&tid[i], NULL, batch_worker, &jobs[i]);
    }
    for (int i = 0; i < threads; i++) pthread_join(tid[i], NULL);
}

/* ---- getSegmentList, batched list form (media-map.js:27-54 per query) ------------ */
/* offsets[Q+1] receives the exclusive prefix of hit counts (offsets[Q] = total hits);
 * hits of query q land at [offsets[q], offsets[q+1]) of sn_out/time_out in fragment
 * order.  Returns the total hit count; hits are only written when total <= cap. */
int64_t oracle_get_segment_list_batch(int32_t n_playlists, const int32_t *playlist_base,
                                      const int64_t *track_off, const uint8_t *track_state,
                                      const double *start, const int32_t *sn, int64_t n_queries,
                                      const int32_t *playlist, const int32_t *level,
                                      const double *begin, const double *duration,
                                      int64_t *offsets, uint8_t *status, int32_t *sn_out,
                                      double *time_out, int64_t cap, int threads)
{
    job_t j;
    memset(&j, 0, sizeof j);
    j.n_playlists = n_playlists; j.playlist_base = playlist_base; j.track_off = track_off;
    j.track_state = track_state; j.start = start; j.sn = sn;
    j.playlist = playlist; j.level = level; j.begin = begin; j.duration = duration;
    j.offsets = offsets; j.status = status; j.sn_out = sn_out; j.time_out = time_out;
    j.cap = cap;
    offsets[0] = 0;
    j.pass = 1;
    run_threads(&j, n_queries, threads);
    for (int64_t q = 0; q < n_queries; q++) offsets[q + 1] += offsets[q];
    int64_t total = offsets[n_queries];
    if (sn_out && time_out && total <= cap) {
        j.pass = 2;
        run_threads(&j, n_queries, threads);
    }
    return total;
}

/* ---- range form helper: (first, count) of the filter on one query ---------------- */
/* Only meaningful when the hits are contiguous (monotone table); `first` is the index
 * of the first kept fragment or -1.  Used to check the CUDA range form. */
void oracle_get_segment_range_batch(int32_t n_playlists, const int32_t *playlist_base,
                                    const int64_t *track_off, const uint8_t *track_state,
                                    const double *start, int64_t n_queries,
                                    const int32_t *playlist, const int32_t *level,
                                    const double *begin, const double *duration,
                                    int32_t *first, int32_t *count, uint8_t *status)
{
    for (int64_t q = 0; q < n_queries; q++) {
        int32_t t = 0;
        int st = resolve_track(n_playlists, playlist_base, track_state,
                               playlist ? playlist[q] : 0, level[q], &t);
        if (status) status[q] = (uint8_t)st;
        first[q] = -1; count[q] = 0;
        if (st != ST_OK) continue;
        int64_t o = track_off[t], n = track_off[t + 1] - o;
        volatile double upper = begin[q] + duration[q];
        double ub = upper;
        int32_t c = 0;
        for (int64_t i = 0; i < n; i++) {
            double s = start[o + i];
            if (begin[q] <= s && s <= ub) { if (!c) first[q] = (int32_t)i; c++; }
        }
        count[q] = c;
    }
}

/* ---- the timed CPU baseline: what the reference does per call -------------------- */
/* One pass per query, hits pushed onto a per-thread growable list exactly like
 * `segmentList.push(new SegmentView(...))` (media-map.js:39-51), then folded into a
 * checksum so the work cannot be elided.  Returns total hits; *checksum_out gets the
 * xor/sum digest.  This is the "C restatement, not Node" baseline of BASELINE.md. */
typedef struct { int32_t sn; int32_t level; int32_t url_id; double time; } view_t;

static void *bench_worker(void *arg)
{
    job_t *j = (job_t *)arg;
    size_t capv = 64;
    view_t *list = (view_t *)malloc(capv * sizeof *list);
    uint64_t cs = 0; int64_t hits = 0;
    for (int64_t q = j->q0; q < j->q1; q++) {
        int32_t t = 0;
        int st = resolve_track(j->n_playlists, j->playlist_base, j->track_state,
                               j->playlist ? j->playlist[q] : 0, j->level[q], &t);
        if (st != ST_OK) continue;
        int64_t o = j->track_off[t], n = j->track_off[t + 1] - o;
        const double *s = j->start + o; const int32_t *snp = j->sn + o;
        double b = j->begin[q];
        volatile double upper = b + j->duration[q];
        double ub = upper;
        size_t len = 0;
        for (int64_t i = 0; i < n; i++) {
            if (b <= s[i] && s[i] <= ub) {
                if (len == capv) { capv *= 2; list = (view_t *)realloc(list, capv * sizeof *list); }
                list[len].sn = snp[i]; list[len].level = j->level[q]; list[len].url_id = 0;
                list[len].time = s[i]; len++;
            }
        }
        for (size_t k = 0; k < len; k++) {
            uint64_t bits; memcpy(&bits, &list[k].time, 8);
            cs += bits * 0x9E3779B97F4A7C15ull + (uint32_t)list[k].sn;
        }
        hits += (int64_t)len;
    }
    free(list);
    j->checksum = cs; j->hits = hits;
    return NULL;
}

int64_t oracle_bench_get_segment_list(int32_t n_playlists, const int32_t *playlist_base,
                                      const int64_t *track_off, const uint8_t *track_state,
                                      const double *start, const int32_t *sn, int64_t n_queries,
                                      const int32_t *playlist, const int32_t *level,
                                      const double *begin, const double *duration, int threads,
                                      uint64_t *checksum_out)
{
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t tid[256];
    job_t jobs[256];
    for (int i = 0; i < threads; i++) {
        memset(&jobs[i], 0, sizeof jobs[i]);
        jobs[i].n_playlists = n_playlists; jobs[i].playlist_base = playlist_base;
        jobs[i].track_off = track_off; jobs[i].track_state = track_state;
        jobs[i].start = start; jobs[i].sn = sn; jobs[i].playlist = playlist;
        jobs[i].level = level; jobs[i].begin = begin; jobs[i].duration = duration;
        jobs[i].q0 = n_queries * i / threads; jobs[i].q1 = n_queries * (i + 1) / threads;
        pthread_create(&tid[i], NULL, bench_worker, &jobs[i]);
    }
    uint64_t cs = 0; int64_t hits = 0;
    for (int i = 0; i < threads; i++) {
        pthread_join(tid[i], NULL);
        cs += jobs[i].checksum; hits += jobs[i].hits;
    }
    if (checksum_out) *checksum_out = cs;
    return hits;
}

/* ---- getSegmentDuration (media-map.js:81-87) -- no reference test, source only --- */
/* Returns the FIRST fragment's duration of the view's level whatever the sn; status 3
 * = empty fragment list (the reference throws 'All segments should have a duration'). */
#define ST_EMPTY 3
void oracle_get_segment_duration_batch(int32_t n_playlists, const int32_t *playlist_base,
                                       const int64_t *track_off, const uint8_t *track_state,
                                       const double *dur0, int64_t n, const int32_t *playlist,
                                       const int32_t *level, double *out, uint8_t *status)
{
    for (int64_t q = 0; q < n; q++) {
        int32_t t = 0;
        int st = resolve_track(n_playlists, playlist_base, track_state,
                               playlist ? playlist[q] : 0, level[q], &t);
        out[q] = NAN;
        if (st == ST_OK) {
            if (track_off[t + 1] == track_off[t]) st = ST_EMPTY;
            else out[q] = dur0[t];
        }
        status[q] = (uint8_t)st;
    }
}

/* ---- getTrackList (media-map.js:60-73): level-major, urlId-minor ----------------- */
int64_t oracle_get_track_list(int32_t n_levels, const int32_t *url_count, int32_t *level_out,
                              int32_t *url_id_out)
{
    int64_t k = 0;
    for (int32_t i = 0; i < n_levels; i++)
        for (int32_t j = 0; j < url_count[i]; j++) {
            if (level_out) { level_out[k] = i; url_id_out[k] = j; }
            k++;
        }
    return k;
}

/* ---- SegmentView wire form (segment-view.js:59-61 / :9-17) ----------------------- */
/* toArrayBuffer: Uint32Array([level, urlId, sn]) -> 12 bytes, ToUint32 on each value,
 * little endian on every platform this runs on; `time` is dropped. */
void oracle_view_pack(int64_t n, const int32_t *level, const int32_t *url_id, const int32_t *sn,
                      uint32_t *wire)
{
    for (int64_t i = 0; i < n; i++) {
        wire[3 * i + 0] = (uint32_t)level[i];
        wire[3 * i + 1] = (uint32_t)url_id[i];
        wire[3 * i + 2] = (uint32_t)sn[i];
    }
}

/* fromArrayBuffer: the first three u32 -> {level, urlId, sn}; values are JS Numbers in
 * [0, 2^32), hence uint32 outputs. */
void oracle_view_unpack(int64_t n, const uint32_t *wire, uint32_t *level, uint32_t *url_id,
                        uint32_t *sn)
{
    for (int64_t i = 0; i < n; i++) {
        level[i] = wire[3 * i + 0];
        url_id[i] = wire[3 * i + 1];
        sn[i] = wire[3 * i + 2];
    }
}

/* isEqual (segment-view.js:33-39 + track-view.js:19-24): sn, level, urlId strictly
 * equal; `time` is not compared. */
void oracle_view_is_equal(int64_t n, const int32_t *al, const int32_t *au, const int32_t *as,
                          const int32_t *bl, const int32_t *bu, const int32_t *bs, uint8_t *out)
{
    for (int64_t i = 0; i < n; i++) out[i] = (as[i] == bs[i] && al[i] == bl[i] && au[i] == bu[i]);
}

/* isInTrack (segment-view.js:45-47): TrackView equality only. */
void oracle_view_is_in_track(int64_t n, const int32_t *al, const int32_t *au, const int32_t *tl,
                             const int32_t *tu, uint8_t *out)
{
    for (int64_t i = 0; i < n; i++) out[i] = (al[i] == tl[i] && au[i] == tu[i]);
}

/* pack -> wire -> parse -> isEqual against the original (test/segment-view.js:21-27).
 * The parsed side holds Numbers in [0, 2^32) while the original holds int32 Numbers, so
 * a negative field does not survive the trip -- compared here as 64-bit integers. */
void oracle_view_roundtrip(int64_t n, const int32_t *level, const int32_t *url_id,
                           const int32_t *sn, uint32_t *wire, uint8_t *equal)
{
    oracle_view_pack(n, level, url_id, sn, wire);
    for (int64_t i = 0; i < n; i++) {
        int64_t pl = wire[3 * i], pu = wire[3 * i + 1], ps = wire[3 * i + 2];
        equal[i] = (ps == (int64_t)sn[i] && pl == (int64_t)level[i] && pu == (int64_t)url_id[i]);
    }
}

/* viewToString (segment-view.js:52-54, track-view.js:11-13): "L{level}U{urlId}S{sn}",
 * plain decimal with a leading '-' for negatives (JS Number->String on an int32).
 * No reference test pins the text -- source only.  offsets[n+1] are byte offsets into
 * `out`; returns the total length.  out may be NULL to size the buffer. */
int64_t oracle_view_format(int64_t n, const int32_t *level, const int32_t *url_id,
                           const int32_t *sn, int with_sn, int64_t *offsets, char *out)
{
    int64_t at = 0;
    char tmp[48];
    for (int64_t i = 0; i < n; i++) {
        int len = with_sn ? snprintf(tmp, sizeof tmp, "L%dU%dS%d", level[i], url_id[i], sn[i])
                          : snprintf(tmp, sizeof tmp, "L%dU%d", level[i], url_id[i]);
        if (offsets) offsets[i] = at;
        if (out) memcpy(out + at, tmp, (size_t)len);
        at += len;
    }
    if (offsets) offsets[n] = at;
    return at;
}

/* Upload-time check the CUDA side must agree with: is the filter a contiguous range?
 * True iff start[] is NaN-free and non-decreasing (hard part 2). */
int oracle_is_monotone(const double *start, int64_t n)
{
    for (int64_t i = 0; i < n; i++) {
        if (start[i] != start[i]) return 0;
        if (i && start[i] < start[i - 1]) return 0;
    }
    return 1;
}

/* ---- SegmentView construction at the loader (p2p-loader-generator.js:145-147) ---------- */
/* new SegmentView({sn: frag.sn, trackView: {level, urlId: level.urlId}, time: frag.start}) for
 * fragment `index` of the level; status 5 = fragments[index] is undefined. */
#define ST_NO_FRAGMENT 5
void oracle_views_from_fragments(int32_t n_playlists, const int32_t *playlist_base,
                                 const int64_t *track_off, const uint8_t *track_state,
                                 const double *start, const int32_t *sn, int64_t n,
                                 const int32_t *playlist, const int32_t *level, const int32_t *index,
                                 int32_t *sn_out, double *time_out, uint8_t *status)
{
    for (int64_t q = 0; q < n; q++) {
        int32_t t = 0;
        int st = resolve_track(n_playlists, playlist_base, track_state, playlist ? playlist[q] : 0,
                               level[q], &t);
        sn_out[q] = 0;
        time_out[q] = NAN;
        if (st == ST_OK) {
            int64_t len = track_off[t + 1] - track_off[t];
            if (index[q] < 0 || index[q] >= len) st = ST_NO_FRAGMENT;
            else {
                sn_out[q] = sn[track_off[t] + index[q]];
                time_out[q] = start[track_off[t] + index[q]];
            }
        }
        status[q] = (uint8_t)st;
    }
}

/* ---- getNextSegmentView: EXTENSION, not in this reference (CHANGELOG.md:89) ------------ */
/* The library's own definition, restated: the fragment following the first fragment of the
 * level whose sn equals the given sn; found = 0 if there is none.  Parity unpinned. */
void oracle_get_next_segment_view(int32_t n_playlists, const int32_t *playlist_base,
                                  const int64_t *track_off, const uint8_t *track_state,
                                  const double *start, const int32_t *sn, int64_t n,
                                  const int32_t *playlist, const int32_t *level, const int32_t *want,
                                  int32_t *sn_out, double *time_out, uint8_t *found, uint8_t *status)
{
    for (int64_t q = 0; q < n; q++) {
        int32_t t = 0;
        int st = resolve_track(n_playlists, playlist_base, track_state, playlist ? playlist[q] : 0,
                               level[q], &t);
        sn_out[q] = 0;
        time_out[q] = NAN;
        found[q] = 0;
        status[q] = (uint8_t)st;
        if (st != ST_OK) continue;
        int64_t o = track_off[t], len = track_off[t + 1] - o;
        for (int64_t i = 0; i < len; i++) {
            if (sn[o + i] == want[q]) {
                if (i + 1 < len) {
                    sn_out[q] = sn[o + i + 1];
                    time_out[q] = start[o + i + 1];
                    found[q] = 1;
                }
                break;
            }
        }
    }
}

/* ---- media playlist text -> fragments (the step before the path) ------- */
/* hls.js 0.5.46 (package.json:34; the dependency is NOT in the reference checkout, so this is
 * restated from that release's published src/loader/playlist-loader.js, parseLevelPlaylist, and
 * the parity it gives is "unpinned": there is no golden vector of the parser to hold it to).
 * One global regular expression is exec'd repeatedly; lastIndex moves past each match.  Only the
 * alternatives that can change fragments[].{sn,start,duration}, `live`, or what a later match can
 * see are restated; TARGETDURATION, DISCONTINUITY and BYTERANGE consume no '#' and set nothing the
 * mapping layer reads, so skipping over them one byte at a time gives the same matches.
 * Decimal conversion is glibc strtod (correctly rounded), the independent check of the device's
 * integer-only converter.  Returns the fragment count; *bad is set when a sequence number leaves
 * int32 (the tables hold sn as int32, media-map.js:46 only copies it). */
static int has_lit(const unsigned char *s, int64_t n, int64_t i, const char *lit)
{
    int64_t k = (int64_t)strlen(lit);
    return i + k <= n && memcmp(s + i, lit, (size_t)k) == 0;
}

static int64_t dot_star_end(const unsigned char *s, int64_t n, int64_t i)
{   /* `.` stops at \n, \r, U+2028, U+2029 (UTF-8 E2 80 A8 / E2 80 A9) */
    while (i < n) {
        if (s[i] == '\n' || s[i] == '\r') break;
        if (s[i] == 0xE2 && i + 2 < n && s[i + 1] == 0x80 && (s[i + 2] == 0xA8 || s[i + 2] == 0xA9)) break;
        i++;
    }
    return i;
}

static int64_t m3u8_parse_one(const unsigned char *s, int64_t n, int32_t *sn_out, double *start_out,
                              double *dur_out, uint8_t *live, int *bad)
{
    int64_t i = 0, count = 0;
    double current_sn = 0.0; /* parseInt gives a Number */
    volatile double total = 0.0;
    *live = 1;
    while (i < n) {
        if (s[i] != '#') { i++; continue; }
        if (has_lit(s, n, i, "#EXT-X-MEDIA-SEQUENCE:")) {
            int64_t j = i + 22;
            if (j < n && s[j] >= '0' && s[j] <= '9') {
                double v = 0.0;
                while (j < n && s[j] >= '0' && s[j] <= '9') { v = v * 10.0 + (s[j] - '0'); j++; }
                if (v > 2147483647.0) { *bad = 1; v = 2147483647.0; }
                current_sn = v;
                i = j;
                continue;
            }
        } else if (has_lit(s, n, i, "#EXT-X-KEY:")) {
            i = dot_star_end(s, n, i + 11);
            continue;
        } else if (has_lit(s, n, i, "#EXT-X-PROGRAM-DATE-TIME:")) {
            i = dot_star_end(s, n, i + 25);
            continue;
        } else if (has_lit(s, n, i, "#EXT-X-ENDLIST")) {
            *live = 0;
            i += 14;
            continue;
        } else if (has_lit(s, n, i, "#EXTINF:")) {
            int64_t j = i + 8, r0 = j;
            while (j < n && ((s[j] >= '0' && s[j] <= '9') || s[j] == '.')) j++;
            if (j > r0) {
                /* parseFloat of the captured run: the StrDecimalLiteral prefix, else NaN */
                char stackbuf[64], *buf = stackbuf;
                int64_t len = j - r0;
                if (len + 1 > (int64_t)sizeof stackbuf) buf = (char *)malloc((size_t)len + 1);
                memcpy(buf, s + r0, (size_t)len);
                buf[len] = 0;
                char *end = buf;
                double d = strtod(buf, &end);
                int is_nan = end == buf;
                if (buf != stackbuf) free(buf);
                while (j < n && s[j] != '\r' && s[j] != '\n') j++;          /* [^\r\n]* */
                int64_t k = j;
                while (k < n && (s[k] == '\r' || s[k] == '\n')) k++;        /* ([\r\n]+[^#|\r\n]+)? */
                if (k > j) {
                    int64_t u = k;
                    while (u < n && s[u] != '#' && s[u] != '|' && s[u] != '\r' && s[u] != '\n') u++;
                    if (u > k) j = u;
                }
                if (!is_nan) {
                    if (current_sn > 2147483647.0) *bad = 1;
                    if (sn_out) {
                        sn_out[count] = (int32_t)(current_sn > 2147483647.0 ? 2147483647.0 : current_sn);
                        start_out[count] = total;
                        dur_out[count] = d;
                    }
                    current_sn += 1.0;
                    total = total + d;
                    count++;
                }
                i = j;
                continue;
            }
        }
        i++;
    }
    return count;
}

typedef struct {
    int32_t t0, t1;
    const uint8_t *state;
    const unsigned char *text;
    const int64_t *text_off;
    int64_t *track_off; /* pass 1: counts at [t+1]; pass 2: offsets */
    uint8_t *live;
    int32_t *sn;
    double *start, *dur;
    int pass, bad;
} m3u8_job_t;

static void *m3u8_worker(void *arg)
{
    m3u8_job_t *j = (m3u8_job_t *)arg;
    for (int32_t t = j->t0; t < j->t1; t++) {
        const unsigned char *s = j->text + j->text_off[t];
        int64_t n = j->text_off[t + 1] - j->text_off[t];
        int parsed = !j->state || j->state[t] == 2;
        if (j->pass == 1) {
            j->live[t] = 1;
            j->track_off[t + 1] = parsed ? m3u8_parse_one(s, n, NULL, NULL, NULL, &j->live[t], &j->bad) : 0;
        } else if (parsed) {
            int64_t o = j->track_off[t];
            m3u8_parse_one(s, n, j->sn + o, j->start + o, j->dur + o, &j->live[t], &j->bad);
        }
    }
    return NULL;
}

/* Returns the total fragment count, or -1 - count when a sequence number left int32.  Call with
 * NULL sn/start/dur to size the columns (track_off and live are filled either way). */
int64_t oracle_parse_m3u8(int32_t n_tracks, const uint8_t *state, const unsigned char *text,
                          const int64_t *text_off, int64_t *track_off, uint8_t *live, int32_t *sn,
                          double *start, double *dur, int threads)
{
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t tid[256];
    m3u8_job_t jobs[256];
    int bad = 0;
    track_off[0] = 0;
    for (int pass = 1; pass <= 2; pass++) {
        if (pass == 2 && !(sn && start && dur)) break;
        for (int i = 0; i < threads; i++) {
            m3u8_job_t *j = &jobs[i];
            j->t0 = (int32_t)((int64_t)n_tracks * i / threads);
            j->t1 = (int32_t)((int64_t)n_tracks * (i + 1) / threads);
            j->state = state; j->text = text; j->text_off = text_off; j->track_off = track_off;
            j->live = live; j->sn = sn; j->start = start; j->dur = dur; j->pass = pass; j->bad = 0;
            if (threads == 1) m3u8_worker(j);
            else pthread_create(&tid[i], NULL, m3u8_worker, j);
        }
        for (int i = 0; i < threads; i++) {
            if (threads > 1) pthread_join(tid[i], NULL);
            bad |= jobs[i].bad;
        }
        if (pass == 1)
            for (int32_t t = 0; t < n_tracks; t++) track_off[t + 1] += track_off[t];
    }
    return bad ? -1 - track_off[n_tracks] : track_off[n_tracks];
}

/* ---- JSON form of a SegmentView (test/segment-view.js:5-11) -------------- */
/* JSON.stringify walks the object the constructor built (segment-view.js:22-26, track-view.js:6-9):
 *   {"sn":S,"trackView":{"level":L,"urlId":U},"time":T}
 * Number::toString (ECMA-262 7.1.12.1): the fewest digits k whose decimal still denotes the same
 * Number, then a layout chosen by the exponent.  Digits here come from glibc -- for k = 1..17 print
 * the correctly rounded k-digit decimal and read it back; at a power of two the interval is lopsided
 * and the k-digit decimal one step up can read back when the rounded one does not -- which is
 * independent of the device's table-driven digit generation.  tests/test_json_host.py holds this
 * against CPython's repr (David Gay's shortest digits) as a third opinion. */
static int js_digits(double x, char *dg, int *n_out)
{   /* x finite, > 0.  digits without trailing zeros into dg; *n_out: value = 0.d1d2.. * 10^n */
    char buf[64], alt[64];
    int p;
    for (p = 1; p <= 17; p++) {
        snprintf(buf, sizeof buf, "%.*e", p - 1, x);
        if (strtod(buf, NULL) == x) break;
        char d2[32];
        int k2 = 0;
        char *e2 = strchr(buf, 'e');
        for (char *c = buf; c < e2; c++) if (*c != '.') d2[k2++] = *c;
        int ex2 = atoi(e2 + 1), i = k2 - 1;
        while (i >= 0 && d2[i] == '9') d2[i--] = '0';
        int at = 0;
        if (i < 0) {
            at = snprintf(alt, sizeof alt, "1e%d", ex2 + 1);
        } else {
            d2[i]++;
            alt[at++] = d2[0];
            alt[at++] = '.';
            for (int j = 1; j < k2; j++) alt[at++] = d2[j];
            if (k2 == 1) alt[at++] = '0';
            at += snprintf(alt + at, sizeof alt - (size_t)at, "e%d", ex2);
        }
        if (strtod(alt, NULL) == x) { memcpy(buf, alt, sizeof buf); break; }
    }
    char *e = strchr(buf, 'e');
    int k = 0;
    for (char *c = buf; c < e; c++) if (*c != '.') dg[k++] = *c;
    while (k > 1 && dg[k - 1] == '0') k--;
    *n_out = atoi(e + 1) + 1;
    return k;
}

/* Number::toString for any double; returns the length (no terminator written beyond it). */
int oracle_js_number_to_string(double x, char *out)
{
    if (x != x) return (int)sprintf(out, "NaN");
    if (x == 0) return (int)sprintf(out, "0");
    int at = 0;
    if (x < 0) { out[at++] = '-'; x = -x; }
    if (x > 1.7976931348623157e308) return at + (int)sprintf(out + at, "Infinity");
    char dg[32];
    int n, k = js_digits(x, dg, &n);
    if (k <= n && n <= 21) {
        for (int i = 0; i < k; i++) out[at++] = dg[i];
        for (int i = k; i < n; i++) out[at++] = '0';
    } else if (0 < n && n <= 21) {
        for (int i = 0; i < n; i++) out[at++] = dg[i];
        out[at++] = '.';
        for (int i = n; i < k; i++) out[at++] = dg[i];
    } else if (-6 < n && n <= 0) {
        out[at++] = '0';
        out[at++] = '.';
        for (int i = 0; i < -n; i++) out[at++] = '0';
        for (int i = 0; i < k; i++) out[at++] = dg[i];
    } else {
        out[at++] = dg[0];
        if (k > 1) {
            out[at++] = '.';
            for (int i = 1; i < k; i++) out[at++] = dg[i];
        }
        at += sprintf(out + at, "e%c%d", n - 1 < 0 ? '-' : '+', n - 1 < 0 ? 1 - n : n - 1);
    }
    out[at] = 0;
    return at;
}

/* Batch JSON.stringify.  time may be NULL (property undefined -> omitted); a non-finite time prints
 * as null.  Call with chars == NULL to size.  Returns the total character count. */
int64_t oracle_view_to_json(int64_t n, const int32_t *level, const int32_t *url_id, const int32_t *sn,
                            const double *time, int32_t *offsets, char *chars)
{
    int64_t at = 0;
    char tmp[160], num[40];
    for (int64_t i = 0; i < n; i++) {
        int len = sprintf(tmp, "{\"sn\":%d,\"trackView\":{\"level\":%d,\"urlId\":%d}", sn[i], level[i], url_id[i]);
        if (time) {
            double t = time[i];
            if (t != t || t - t != 0) len += sprintf(tmp + len, ",\"time\":null");
            else { oracle_js_number_to_string(t, num); len += sprintf(tmp + len, ",\"time\":%s", num); }
        }
        tmp[len++] = '}';
        offsets[i] = (int32_t)at;
        if (chars) memcpy(chars + at, tmp, (size_t)len);
        at += len;
    }
    offsets[n] = (int32_t)at;
    return at;
}

/* new SegmentView(JSON.parse(text)) for texts in the shape above (keys in constructor order, each
 * optional, no white space).  flags: 1 sn, 2 trackView, 4 level, 8 urlId, 16 time, 32 time is null.
 * status: 0 ok, 1 malformed, 3 an int column that is not an int32.  (Status 2, "number outside the
 * exact converter", is the device's own limit: strtod has none, so the tests keep such inputs apart.) */
static int json_num(const char **pp, const char *end, double *out)
{   /* RFC 8259 number at *pp; advances; 0 if not a number */
    const char *p = *pp, *q = p;
    if (q < end && *q == '-') q++;
    if (q >= end || *q < '0' || *q > '9') return 0;
    if (*q == '0') { q++; if (q < end && *q >= '0' && *q <= '9') return 0; }
    else while (q < end && *q >= '0' && *q <= '9') q++;
    if (q < end && *q == '.') {
        q++;
        if (q >= end || *q < '0' || *q > '9') return 0;
        while (q < end && *q >= '0' && *q <= '9') q++;
    }
    if (q < end && (*q == 'e' || *q == 'E')) {
        q++;
        if (q < end && (*q == '+' || *q == '-')) q++;
        if (q >= end || *q < '0' || *q > '9') return 0;
        while (q < end && *q >= '0' && *q <= '9') q++;
    }
    char buf[400];
    size_t len = (size_t)(q - p);
    if (len >= sizeof buf) return 0;
    memcpy(buf, p, len);
    buf[len] = 0;
    *out = strtod(buf, NULL);
    *pp = q;
    return 1;
}

static int json_lit(const char **pp, const char *end, const char *lit)
{
    size_t k = strlen(lit);
    if ((size_t)(end - *pp) < k || memcmp(*pp, lit, k) != 0) return 0;
    *pp += k;
    return 1;
}

static int json_int(const char **pp, const char *end, int32_t *out)
{
    double x;
    if (!json_num(pp, end, &x)) return 1;
    if (!(x >= -2147483648.0 && x <= 2147483647.0) || x != (double)(int32_t)x) return 3;
    *out = (int32_t)x;
    return 0;
}

void oracle_view_from_json(int64_t n, const char *chars, const int32_t *offsets, int32_t *level,
                           int32_t *url_id, int32_t *sn, double *time, uint8_t *flags, uint8_t *status)
{
    for (int64_t i = 0; i < n; i++) {
        const char *p = chars + offsets[i], *end = chars + offsets[i + 1];
        int st = 0, first = 1;
        unsigned f = 0;
        sn[i] = level[i] = url_id[i] = 0;
        time[i] = NAN;
        do {
            if (!json_lit(&p, end, "{")) { st = 1; break; }
            if (json_lit(&p, end, "\"sn\":")) {
                if ((st = json_int(&p, end, &sn[i]))) break;
                f |= 1; first = 0;
            }
            if (json_lit(&p, end, first ? "\"trackView\":{" : ",\"trackView\":{")) {
                int inner_first = 1;
                f |= 2; first = 0;
                if (json_lit(&p, end, "\"level\":")) {
                    if ((st = json_int(&p, end, &level[i]))) break;
                    f |= 4; inner_first = 0;
                }
                if (json_lit(&p, end, inner_first ? "\"urlId\":" : ",\"urlId\":")) {
                    if ((st = json_int(&p, end, &url_id[i]))) break;
                    f |= 8;
                }
                if (!json_lit(&p, end, "}")) { st = 1; break; }
            }
            if (json_lit(&p, end, first ? "\"time\":" : ",\"time\":")) {
                f |= 16;
                if (json_lit(&p, end, "null")) f |= 32;
                else if (!json_num(&p, end, &time[i])) { st = 1; break; }
            }
            if (!json_lit(&p, end, "}") || p != end) st = 1;
        } while (0);
        if (st) { /* columns of a rejected text hold what was read before the failure */ }
        flags[i] = (uint8_t)f;
        status[i] = (uint8_t)st;
    }
}

// common.cuh -- shared declarations of libhlsmap (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "hlsmap.h"

namespace hlsmap {

// One fragment table = one (playlist, level).  16 bytes so a query fetches it with a
// single LDG.128.  flags: bits 0-1 = HLSMAP_TRACK_*, bit 2 = start column is monotone
// non-decreasing and NaN-free (established on the GPU at upload).
struct __align__(16) TrackDesc {
    long long off;   // first fragment in the start[] / sn[] columns
    int len;         // fragments in this track
    unsigned flags;
};
// Hint grid of one track (SMALL maps only), 32 bytes = two LDG.128.
struct __align__(16) TrackHint {
    double t0;     // start[0]
    double inv;    // ncell / (start[len-1] - start[0]), or 0 when degenerate
    int grid_off;  // first entry of this track in the u16 grid (ncell + 1 entries)
    int ncell;
    int pad[2];
};
constexpr unsigned kTrackStateMask = 3u;
constexpr unsigned kTrackMonotone = 4u;
constexpr unsigned kTrackSnContig = 8u;   // bit 3: sn[i] == sn[0] + i for the whole track (proven at upload)

// Per-context scratch header for the single-pass scan kernels.  A context owns TWO scratch
// sets (header + status words + result) used by alternate launches (set = epoch & 1), so that
// a batch may start while the previous one is still draining; the launches on a set are numbered
// and take it in that order (`turn`: how many have finished with it; passed on by the last CTA of
// a launch to exit, which `exited` counts), and ticket[] is double buffered: the host tells
// each launch which of the two counters to use (they alternate per set), and a launch's last
// tile zeroes the OTHER one, so late ticket fetches of a launch never see the next one's reset.
struct ScanCounters {
    unsigned int ticket[2];
    unsigned int done;
    unsigned int turn;        // launches that have finished with this set (the next one to use it has seq == turn)
    unsigned long long n_unparsed;
    unsigned long long n_missing;
    unsigned int exited;      // CTAs of the owning launch that have left
    unsigned int pad;
};

// What the last tile of a batch leaves behind for the host.
struct DeviceResult {
    long long total_hits;
    long long n_unparsed;
    long long n_missing;
    long long overflow;  // 1 if the running hit count passed INT32_MAX
};

// Tile status word of the decoupled look-back: [epoch:22][state:2][value:40].
constexpr unsigned long long kStateAggregate = 1ull;
constexpr unsigned long long kStatePrefix = 2ull;
constexpr int kValueBits = 40;
constexpr unsigned kEpochMask = (1u << 22) - 1u;

__host__ __device__ inline unsigned long long make_status(unsigned epoch, unsigned long long state,
                                                          unsigned long long value)
{
    return ((unsigned long long)(epoch & kEpochMask) << 42) | (state << kValueBits) |
           (value & ((1ull << kValueBits) - 1ull));
}

// One list-form batch inside a launch of the search kernel.  A launch resolves up to kMaxFused
// independent batches (hlsmap_get_segment_list_dev_batches under hlsmap_ctx_set_batch_overlap): their
// tiles are numbered consecutively, tickets hand them out in that order, and a CTA stages the tables
// once for all of them -- n batches cost one ramp and one tail instead of n.
struct BatchDesc {
    const int *q_playlist;  // nullable
    const int *q_level;
    const double *q_begin;
    const double *q_duration;
    int *offsets;           // [nq+1]
    int *sn_out;            // nullable (count only)
    double *time_out;
    unsigned char *status;  // nullable
    long long nq;
    long long capacity;
    unsigned long long base;   // hits produced by earlier chunks of the same call (0 normally)
    int tile_first;            // the launch-wide number of the batch's first tile
    int n_tiles;
    int vec_ok;                // query/offset pointers are 16-byte aligned
    int pad;
};
constexpr int kMaxFused = 32;

struct ResolveParams {
    // tables
    const TrackDesc *desc;
    const int *playlist_base;
    int n_playlists;
    const double *start;
    const int *sn;
    long long n_frag_padded;  // elements readable in start[] (multiple of 4)
    const TrackHint *hint;        // SMALL maps only
    const unsigned short *grid;   // SMALL maps only
    int grid_padded;              // u16 entries in grid (multiple of 8)
    int n_tracks_padded;          // SMALL: descriptors staged in shared memory (>= 1)
    int n_base_padded;            // SMALL: playlist_base ints staged (multiple of 4)
    int small;                    // whole map staged in shared memory + hint grid
    int uniform_levels;           // > 0: every playlist has exactly this many levels (track = playlist * that + level)
    int sn_contig;                // every parsed track's sn column is contiguous (general kernel computes sn)
    int ipt;                      // queries per thread of the SMALL kernel (2 or 4)
    // queries
    long long nq;
    const int *q_playlist;  // nullable
    const int *q_level;
    const double *q_begin;
    const double *q_duration;
    // outputs
    int *offsets;           // [nq+1]
    int *sn_out;            // nullable (count only)
    double *time_out;
    long long capacity;
    unsigned char *status;  // nullable
    // scratch
    unsigned long long *tile_status;
    ScanCounters *ctr;
    DeviceResult *result;
    unsigned epoch;
    unsigned tslot;         // the launch's number on its scratch set (0, 1, 2, ...); its low bit picks the ticket counter
    int n_tiles;
    // staging
    int stage_elems;  // capacity of the shared-memory copy of start[], in doubles
    int stage_all;    // the whole column fits: stage once per CTA
    int vec_ok;       // query/offset pointers are 16-byte aligned
    int tile_q;       // queries per tile (<= 256 * ipt, multiple of 4)
    unsigned long long base;  // hits produced by earlier chunks of the same call (0 normally)
    unsigned long long *trace;  // diagnostics: per-tile phase timestamps (HLSMAP_TRACE=file); NULL normally
};
// What a launch of the search kernel takes: the launch-wide fields above (n_tiles = the launch's total)
// and its batches.  NB = 1: one batch, the usual call -- a small parameter block, batch index 0 at
// compile time; NB = kMaxFused: the fused launches of a batches call.
template <int NB>
struct SearchParams : ResolveParams {
    int n_batches;
    BatchDesc fb[NB];
};
static_assert(sizeof(SearchParams<kMaxFused>) <= 4096, "kernel parameters are limited to 4 KB");

struct RangeParams {
    const TrackDesc *desc;
    const int *playlist_base;
    int n_playlists;
    const double *start;
    long long nq;
    const int *q_playlist;
    const int *q_level;
    const double *q_begin;
    const double *q_duration;
    int *first;
    int *count;
    unsigned char *status;
};

// ---- playlist ingest (ingest.cu) -----------------------------------------------------------
constexpr int kIngestThreads = 128;
constexpr int kIngestTileBytes = kIngestThreads * 128;
struct IngestCounters {
    unsigned long long n_fragments;
    unsigned long long n_tags;
    unsigned long long n_unsupported;
};
struct IngestParams {
    const unsigned char *text;   // all playlists, concatenated
    long long n_bytes;
    const long long *text_off;   // [n_tracks + 1]
    int n_tracks;
    const unsigned char *state;  // nullable: every track is parsed
    const int *tile_tlo;         // [n_tiles + 1] track holding the tile's first byte
    const int *tile_bfirst;      // [n_tiles + 1] first track starting at/after the tile's first byte
    double *dur;                 // durations in playlist order
    long long capacity;
    long long *track_off;        // [n_tracks + 1]
    unsigned char *live;         // [n_tracks], preset to 1; ENDLIST clears it
    long long *tag_pos;          // live #EXT-X-MEDIA-SEQUENCE tags: text position,
    long long *tag_frag;         //   fragments before the tag,
    int *tag_val;                //   and the number
    long long tag_capacity;
    // No scan inside the parse: every 4 KB span of text leaves its fragments and tags in a slot
    // of fixed size and its two counts; m3u8_compact_kernel moves them to their final places.
    double *stage_dur;           // [n_spans * kIngestSpanFrags]; nullptr = count only
    unsigned long long *stage_tag;   // [n_spans * kIngestSpanTags]: value | fragments before << 32 | position in span << 48
    unsigned *span_cnt;          // [n_spans]: fragments | tags << 16
    unsigned long long *block_sum;   // [n_blocks]: per kIngestScanBlock spans, fragments | tags << 40 (zeroed)
    unsigned long long *block_base;  // [n_blocks + 1]: exclusive scan of block_sum
    unsigned long long *span_base;   // [n_spans]: fragments | tags << 40 before each span
    int n_tiles, n_spans, n_blocks;
    int vec_ok;                  // text is 16-byte aligned
    IngestCounters *out;
};
constexpr int kIngestSpanBytes = 4096;
constexpr int kIngestSpanFrags = 456;   // a fragment takes at least "#EXTINF:0" and a line break
constexpr int kIngestSpanTags = 180;    // "#EXT-X-MEDIA-SEQUENCE:0" is 23 bytes
constexpr int kIngestScanBlock = 1024;  // spans per block of the two-level scan
struct AssembleParams {
    int n_tracks;
    const unsigned char *state;
    const long long *text_off;
    const long long *track_off;
    const double *dur;
    const long long *tag_pos, *tag_frag;
    const int *tag_val;
    long long capacity, tag_capacity;   // of the start/sn/dur columns and of the tag list
    double *start;
    int *sn;
    double *dur0;
    IngestCounters *out;
};
cudaError_t launch_ingest_tile_index(const long long *text_off, int n_tracks, int n_tiles, int *tile_tlo,
                                     int *tile_bfirst, cudaStream_t s);
cudaError_t launch_ingest_scan(const IngestParams &p, cudaStream_t s);
cudaError_t launch_ingest_compact(const IngestParams &p, cudaStream_t s);   // block scan, compaction, track offsets
cudaError_t launch_ingest_assemble(const AssembleParams &p, bool warp_per_track, cudaStream_t s);

// ---- JSON form of views (json.cu) -----------------------------------------------------------
constexpr int kJsonTileViews = 256;
struct JsonWriteParams {
    long long n;
    const int *level, *url_id, *sn;
    const double *time;          // nullable: the property is undefined and left out
    int *offsets;
    char *chars;
    long long capacity;
    unsigned long long *stage_digits;   // [n] shortest digits of time
    unsigned *stage_meta;               // [n] exp10 | ndigits << 16 | kind << 21 | negative << 23 | length << 24
    unsigned long long *block_sum;      // [n_tiles] characters per 256-view tile
    unsigned long long *block_base;     // [n_tiles] characters before each tile
    DeviceResult *result;
    int n_tiles;
};
struct JsonReadParams {
    long long n;
    int vec_ok;                  // chars is 16-byte aligned
    const char *chars;
    const int *offsets;
    int *level, *url_id, *sn;
    double *time;
    unsigned char *flags, *status;
};
int view_to_json_max_grid();
cudaError_t launch_view_to_json(const JsonWriteParams &p, int grid, cudaStream_t s);
cudaError_t launch_view_from_json(const JsonReadParams &p, cudaStream_t s);

// launchers (resolve.cu)
cudaError_t launch_resolve_search(const ResolveParams &p, int grid, size_t smem, cudaStream_t s,
                                  bool overlap_next = false);
cudaError_t launch_resolve_search_fused(const ResolveParams &p, const BatchDesc *b, int n, int grid, size_t smem,
                                        cudaStream_t s, bool overlap_next);
cudaError_t launch_resolve_filter(const ResolveParams &p, int grid, cudaStream_t s);
cudaError_t launch_range(const RangeParams &p, cudaStream_t s);
cudaError_t launch_range_small(const RangeParams &p, const TrackHint *hint, const unsigned short *grid,
                               long long n_frag_padded, int grid_padded, int n_tracks_padded,
                               int n_base_padded, int vec_ok, cudaStream_t s);
cudaError_t launch_monotone_check(const double *start, const int *sn, TrackDesc *desc, int n_tracks,
                                  cudaStream_t s);
cudaError_t launch_segment_duration(const TrackDesc *desc, const int *playlist_base,
                                    int n_playlists, const double *dur0, long long n,
                                    const int *q_playlist, const int *q_level, double *out,
                                    unsigned char *status, cudaStream_t s);
cudaError_t launch_slide_tracks(double *start, int *sn, double *dur0, TrackDesc *desc, long long n,
                                const int *track, const int *drop, const long long *append_off,
                                const double *new_start, const int *new_sn, const double *new_dur0,
                                cudaStream_t s);
cudaError_t launch_views_from_fragments(const TrackDesc *desc, const int *playlist_base, int n_playlists,
                                        const double *start, const int *sn, long long n,
                                        const int *q_playlist, const int *q_level, const int *q_index,
                                        int *out_sn, double *out_time, unsigned char *status,
                                        cudaStream_t s);
cudaError_t launch_next_segment_view(const TrackDesc *desc, const int *playlist_base, int n_playlists,
                                     const double *start, const int *sn, long long n,
                                     const int *q_playlist, const int *q_level, const int *q_sn,
                                     int *out_sn, double *out_time, unsigned char *found,
                                     unsigned char *status, cudaStream_t s);
int resolve_search_max_grid(size_t smem, bool small, int ipt);
int resolve_search_tile(bool small, int ipt);
cudaError_t launch_build_hint(const double *start, const TrackDesc *desc, TrackHint *hint,
                              unsigned short *grid, int n_tracks, cudaStream_t s);
int resolve_filter_max_grid();
constexpr int kFilterTileQueries = 256;   // 8 warps x 32 queries

// launchers (views.cu)
struct FormatParams {
    long long n;
    const int *level, *url_id, *sn;
    int with_sn;
    int *offsets;
    char *chars;
    long long capacity;
    unsigned long long *tile_status;
    ScanCounters *ctr;
    DeviceResult *result;
    unsigned epoch;
    unsigned tslot;
    int n_tiles;
    unsigned long long *block_sum, *block_base;   // [n_tiles] each: the three-launch form
};
constexpr int kFormatTileViews = 1024;
cudaError_t launch_view_pack(long long n, const int *level, const int *url_id, const int *sn,
                             unsigned *wire, cudaStream_t s);
cudaError_t launch_view_unpack(long long n, const unsigned *wire, unsigned *level,
                               unsigned *url_id, unsigned *sn, cudaStream_t s);
cudaError_t launch_view_is_equal(long long n, const int *al, const int *au, const int *as,
                                 const int *bl, const int *bu, const int *bs, unsigned char *out,
                                 cudaStream_t s);
cudaError_t launch_view_is_in_track(long long n, const int *al, const int *au, const int *tl,
                                    const int *tu, unsigned char *out, cudaStream_t s);
cudaError_t launch_view_roundtrip(long long n, const int *level, const int *url_id, const int *sn,
                                  unsigned *wire, unsigned char *equal, cudaStream_t s);
cudaError_t launch_view_format(const FormatParams &p, int grid, cudaStream_t s);
cudaError_t launch_view_format_split(const FormatParams &p, cudaStream_t s);
int view_format_max_grid();

}  // namespace hlsmap

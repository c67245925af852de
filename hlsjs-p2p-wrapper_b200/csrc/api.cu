// api.cu -- the extern "C" boundary of libhlsmap: contexts, table upload, and the
// host-pointer / device-pointer entry points declared in include/hlsmap.h.
//
// No CPU fallback lives here: every entry point either runs the CUDA kernels of
// resolve.cu / views.cu or fails with a message.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <new>
#include <string>
#include <vector>

#include "common.cuh"

using namespace hlsmap;

#ifndef HLSMAP_TRACE_BUILD
#define HLSMAP_TRACE_BUILD 0
#endif

namespace {

thread_local std::string g_err;

int fail(int code, const std::string &msg)
{
    g_err = msg;
    return code;
}

#define CK(call)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess)                                                           \
            return fail(HLSMAP_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

constexpr int kSlots = 12;

}  // namespace

struct hlsmap_ctx {
    int device = 0;
    int sm_count = 0;
    int max_smem_optin = 0;
    cudaStream_t stream = nullptr;      // compute + copies of the host-pointer variants
    cudaStream_t s_in = nullptr;        // host pipeline: query uploads
    cudaStream_t s_out = nullptr;       // host pipeline: result downloads
    cudaEvent_t ev_in[2] = {nullptr, nullptr};
    ScanCounters *ctr = nullptr;        // scan scratch headers, one per scratch set (2)
    DeviceResult *d_result = nullptr;   // results, one per scratch set (2)
    bool overlap = false;               // consecutive search batches may overlap (PDL)
    int64_t host_chunk = 262144;        // largest chunk of that pipeline, in queries (HLSMAP_HOST_CHUNK at creation, experiments)
    bool host_pipeline = true;          // chunked upload/compute/download of large host-pointer calls
    void *begin_event = nullptr;        // a batches call's opening cudaEvent_t, recorded right before its first launch
    bool env_no_fuse = false;
    bool no_fuse = false;               // HLSMAP_NO_FUSE: a batches call issues one launch per batch even when it may fuse them
                                        // (HLSMAP_NO_PIPELINE in the environment at context creation turns it off)
    unsigned set_launches[2] = {0, 0};  // kernels launched on each scratch set (ticket slot parity)
    DeviceResult *h_result = nullptr;   // pinned
    unsigned long long *tile_status = nullptr;   // 2 x tile_cap status words (set = epoch & 1)
    int64_t tile_cap = 0;
    unsigned epoch = 0;
    int last_launches = 0;
    int last_path = 0;
    bool have_last = false;
    void *slot[kSlots] = {};            // grow-only device staging of the host variants
    int64_t slot_bytes[kSlots] = {};
};

struct hlsmap_map {
    hlsmap_ctx *ctx = nullptr;
    int32_t n_playlists = 0, n_tracks = 0;
    int64_t n_frag = 0, n_frag_padded = 0;
    TrackDesc *d_desc = nullptr;
    int *d_playlist_base = nullptr;
    double *d_start = nullptr;
    int *d_sn = nullptr;
    double *d_dur0 = nullptr;
    TrackHint *d_hint = nullptr;        // SMALL maps: per-track hint-grid parameters
    unsigned short *d_grid = nullptr;   // SMALL maps: the hint grid itself
    int grid_padded = 0;
    int n_tracks_padded = 1, n_base_padded = 4;
    std::vector<int32_t> h_playlist_base;
    std::vector<int32_t> h_url_count;
    std::vector<uint8_t> h_monotone;
    std::vector<uint8_t> h_state;       // HLSMAP_TRACK_* per track
    std::vector<int32_t> h_len, h_cap;  // current and reserved fragments per track (live updates)
    hlsmap_map_info info{};
    int path = HLSMAP_PATH_AUTO;
    int stage_elems = 0;
    size_t stage_smem = 0;
    int grid_search = 0, grid_filter = 0;
    int ipt = 2;                        // queries per thread of the search kernel (tiles of 256 * ipt)
    int grid_search_wide = 0;           // SMALL maps: grid of the 1024-query-tile variant (0 = not offered)
    int uniform_levels = 0;             // > 0: every playlist has exactly this many levels
    bool sn_contig = false;             // every parsed track numbers its fragments contiguously (proven on the GPU)
};

namespace {

int use_ctx(hlsmap_ctx *ctx)
{
    if (!ctx) return fail(HLSMAP_ERR_INVALID, "null context");
    CK(cudaSetDevice(ctx->device));
    return HLSMAP_OK;
}

// `user` = the stream of a device-pointer call: work queued there by an earlier asynchronous call
// may still be using the slot that is about to be replaced.
int ensure_slot(hlsmap_ctx *ctx, int i, int64_t bytes, void **out, cudaStream_t user = nullptr)
{
    if (bytes < 16) bytes = 16;
    if (ctx->slot_bytes[i] < bytes) {
        if (ctx->slot[i]) {
            if (user && user != ctx->stream) CK(cudaStreamSynchronize(user));
            CK(cudaStreamSynchronize(ctx->stream));
            CK(cudaFree(ctx->slot[i]));
            ctx->slot[i] = nullptr;
            ctx->slot_bytes[i] = 0;
        }
        const int64_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&ctx->slot[i], (size_t)want);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(HLSMAP_ERR_NOMEM, "cudaMalloc of " + std::to_string(want) + " bytes failed");
        }
        ctx->slot_bytes[i] = want;
    }
    *out = ctx->slot[i];
    return HLSMAP_OK;
}

// The opening event of a timed batches call goes onto the stream immediately before the call's first
// launch (after the host-side preparation, which would otherwise count as device time on an idle stream).
inline void mark_begin(hlsmap_ctx *ctx, cudaStream_t s)
{
    if (ctx->begin_event) {
        cudaEventRecord((cudaEvent_t)ctx->begin_event, s);
        ctx->begin_event = nullptr;
    }
}

// Room for the status words of n_tiles tiles in each of the two scratch sets.  Grows by reallocating
// (stream synchronisation, cudaFree, cudaMalloc, a clearing memset): callers that time what follows
// reserve first.  The first allocation holds 65536 tiles (2 MiB), which most callers never outgrow.
int reserve_tiles(hlsmap_ctx *ctx, int64_t n_tiles, cudaStream_t s)
{
    if (n_tiles <= ctx->tile_cap) return HLSMAP_OK;
    CK(cudaStreamSynchronize(ctx->stream));
    if (s != ctx->stream) CK(cudaStreamSynchronize(s));
    if (ctx->tile_status) CK(cudaFree(ctx->tile_status));
    ctx->tile_status = nullptr;
    int64_t cap = n_tiles + n_tiles / 2 + 1024;
    if (cap < 65536) cap = 65536;
    cudaError_t e = cudaMalloc(&ctx->tile_status, (size_t)cap * 16);
    if (e != cudaSuccess) {
        cudaGetLastError();
        ctx->tile_cap = 0;
        return fail(HLSMAP_ERR_NOMEM, "cudaMalloc of tile status failed");
    }
    ctx->tile_cap = cap;
    CK(cudaMemsetAsync(ctx->tile_status, 0, (size_t)cap * 16, s));
    return HLSMAP_OK;
}

// Scratch for a scan batch of n_tiles tiles; bumps the launch epoch.  Launches alternate
// between two scratch sets (epoch & 1) so that a batch may start while its predecessor drains.
int prepare_scan(hlsmap_ctx *ctx, int64_t n_tiles, cudaStream_t s)
{
    {
        const int r = reserve_tiles(ctx, n_tiles, s);
        if (r) return r;
    }
    ctx->epoch++;
    if (ctx->epoch > kEpochMask) {  // epoch 0 means "never written": wipe and restart
        CK(cudaStreamSynchronize(s));
        CK(cudaMemsetAsync(ctx->tile_status, 0, (size_t)ctx->tile_cap * 16, s));
        CK(cudaMemsetAsync(ctx->ctr, 0, 2 * sizeof(ScanCounters), s));
        ctx->set_launches[0] = ctx->set_launches[1] = 0;
        ctx->epoch = 1;
    }
    return HLSMAP_OK;
}

inline ScanCounters *ctr_of(hlsmap_ctx *ctx) { return ctx->ctr + (ctx->epoch & 1u); }
inline DeviceResult *result_of(hlsmap_ctx *ctx) { return ctx->d_result + (ctx->epoch & 1u); }
inline unsigned long long *status_of(hlsmap_ctx *ctx)
{
    return ctx->tile_status + (ctx->epoch & 1u) * ctx->tile_cap;
}

int fetch_result(hlsmap_ctx *ctx, cudaStream_t s, hlsmap_result *out)
{
    CK(cudaMemcpyAsync(ctx->h_result, result_of(ctx), sizeof(DeviceResult), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    out->total_hits = ctx->h_result->total_hits;
    out->n_unparsed = ctx->h_result->n_unparsed;
    out->n_missing = ctx->h_result->n_missing;
    out->kernel_launches = ctx->last_launches;
    out->path = ctx->last_path;
    if (ctx->h_result->overflow)
        return fail(HLSMAP_ERR_OVERFLOW, "batch produced more than INT32_MAX hits; split it");
    return HLSMAP_OK;
}

int pick_path(const hlsmap_map *m)
{
    if (m->path == HLSMAP_PATH_FILTER) return HLSMAP_PATH_FILTER;
    if (m->path == HLSMAP_PATH_SEARCH) return HLSMAP_PATH_SEARCH;
    return m->info.all_monotone ? HLSMAP_PATH_SEARCH : HLSMAP_PATH_FILTER;
}

inline bool al16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Tiling of a batch: full tiles of the kernel's 256*ipt queries.  (Measured on B200 and lost, both
// against uniform 512-query tiles on the 10^6-query workload: tile size trimmed so that every CTA
// gets a whole number of tiles, 61 us vs 57 us; the last 15-40 % of a lone batch in tiles of half
// or a quarter the size, 52-67 us vs 51.7 -- per-tile fixed costs outweigh the idle tail round.)
struct TilePlan {
    int tile_q = 0;
    int64_t n_tiles = 0;
};
TilePlan plan_tiles(const hlsmap_map *map, int path, int64_t nq, int ipt = 0)
{
    TilePlan t;
    t.tile_q = path == HLSMAP_PATH_SEARCH ? resolve_search_tile(map->info.stage_all != 0, ipt ? ipt : map->ipt)
                                          : kFilterTileQueries;
    t.n_tiles = (nq + t.tile_q - 1) / t.tile_q;
    return t;
}

}  // namespace

extern "C" {

const char *hlsmap_last_error(void) { return g_err.c_str(); }
int hlsmap_version(void) { return 200; }

int hlsmap_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(HLSMAP_ERR_CUDA, std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e));
    }
    return n;
}

int hlsmap_ctx_create(int device, hlsmap_ctx **out)
{
    if (!out) return fail(HLSMAP_ERR_INVALID, "null out pointer");
    *out = nullptr;
    int n = hlsmap_device_count();
    if (n < 0) return n;
    if (n == 0) return fail(HLSMAP_ERR_CUDA, "no CUDA device: libhlsmap has no CPU path");
    if (device < 0 || device >= n) return fail(HLSMAP_ERR_INVALID, "device ordinal out of range");
    CK(cudaSetDevice(device));
    hlsmap_ctx *c = new (std::nothrow) hlsmap_ctx();
    if (!c) return fail(HLSMAP_ERR_NOMEM, "out of host memory");
    c->device = device;
    c->host_pipeline = getenv("HLSMAP_NO_PIPELINE") == nullptr;   // read once, here
    c->env_no_fuse = getenv("HLSMAP_NO_FUSE") != nullptr;         // experiments: one launch per batch of a batches call
    c->no_fuse = c->env_no_fuse;
    if (const char *e = getenv("HLSMAP_HOST_CHUNK")) {
        const long long v = atoll(e);
        if (v >= 1024) c->host_chunk = (v + 1023) / 1024 * 1024;
    }
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        delete c;
        return fail(HLSMAP_ERR_CUDA, "libhlsmap is built for sm_100a (Blackwell B200) only");
    }
    c->sm_count = prop.multiProcessorCount;
    c->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
    CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&c->s_in, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&c->s_out, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&c->ev_in[0], cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&c->ev_in[1], cudaEventDisableTiming));
    CK(cudaMalloc(&c->ctr, 2 * sizeof(ScanCounters)));
    CK(cudaMalloc(&c->d_result, 2 * sizeof(DeviceResult)));
    CK(cudaMemset(c->ctr, 0, 2 * sizeof(ScanCounters)));
    CK(cudaMemset(c->d_result, 0, 2 * sizeof(DeviceResult)));
    CK(cudaHostAlloc(&c->h_result, sizeof(DeviceResult), cudaHostAllocDefault));
    memset(c->h_result, 0, sizeof(DeviceResult));
    *out = c;
    return HLSMAP_OK;
}

void hlsmap_ctx_destroy(hlsmap_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (int i = 0; i < kSlots; i++)
        if (c->slot[i]) cudaFree(c->slot[i]);
    if (c->tile_status) cudaFree(c->tile_status);
    if (c->ctr) cudaFree(c->ctr);
    if (c->d_result) cudaFree(c->d_result);
    if (c->h_result) cudaFreeHost(c->h_result);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->s_in) cudaStreamDestroy(c->s_in);
    if (c->s_out) cudaStreamDestroy(c->s_out);
    if (c->ev_in[0]) cudaEventDestroy(c->ev_in[0]);
    if (c->ev_in[1]) cudaEventDestroy(c->ev_in[1]);
    delete c;
}

int hlsmap_ctx_device(const hlsmap_ctx *ctx) { return ctx ? ctx->device : -1; }

int hlsmap_ctx_set_batch_overlap(hlsmap_ctx *ctx, int enable)
{
    if (!ctx) return fail(HLSMAP_ERR_INVALID, "null context");
    ctx->overlap = enable != 0;
    ctx->no_fuse = ctx->env_no_fuse || enable == 2;      // 2: overlapping launches, one per batch
    return HLSMAP_OK;
}

int hlsmap_ctx_synchronize(hlsmap_ctx *ctx)
{
    int r = use_ctx(ctx);
    if (r) return r;
    CK(cudaStreamSynchronize(ctx->stream));
    return HLSMAP_OK;
}

int hlsmap_host_alloc(hlsmap_ctx *ctx, int64_t bytes, void **out)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (!out || bytes < 0) return fail(HLSMAP_ERR_INVALID, "bad host_alloc arguments");
    CK(cudaHostAlloc(out, (size_t)(bytes > 0 ? bytes : 1), cudaHostAllocPortable));
    return HLSMAP_OK;
}

int hlsmap_host_free(hlsmap_ctx *ctx, void *p)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (p) CK(cudaFreeHost(p));
    return HLSMAP_OK;
}

// ---------------------------------------------------------------------------------------
//  table upload
// ---------------------------------------------------------------------------------------
// `start` / `sn` may live on the host or on the device (the copies below let the runtime tell);
// `duration` is a host column or NULL; `dev_dur0`, when given, is a device array [n_tracks] that
// already holds every track's first duration (tables parsed on the device).
static int map_create_impl(hlsmap_ctx *ctx, int32_t n_playlists, const int32_t *playlist_track_base,
                           int32_t n_tracks, const int64_t *track_offset, const uint8_t *track_state,
                           const int32_t *url_count, const double *start, const int32_t *sn,
                           const double *duration, const double *dev_dur0, hlsmap_map **out)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (!out) return fail(HLSMAP_ERR_INVALID, "null out pointer");
    *out = nullptr;
    if (n_playlists < 0 || n_tracks < 0 || !playlist_track_base || !track_offset ||
        (n_tracks > 0 && !track_state))
        return fail(HLSMAP_ERR_INVALID, "map_create: missing table arrays");
    if (playlist_track_base[0] != 0 || playlist_track_base[n_playlists] != n_tracks)
        return fail(HLSMAP_ERR_INVALID, "map_create: playlist_track_base must span [0, n_tracks]");
    for (int32_t p = 0; p < n_playlists; p++)
        if (playlist_track_base[p + 1] < playlist_track_base[p])
            return fail(HLSMAP_ERR_INVALID, "map_create: playlist_track_base not monotone");
    if (track_offset[0] != 0) return fail(HLSMAP_ERR_INVALID, "map_create: track_offset[0] != 0");
    for (int32_t t = 0; t < n_tracks; t++) {
        const int64_t len = track_offset[t + 1] - track_offset[t];
        if (len < 0 || len > 0x7fffffff)
            return fail(HLSMAP_ERR_INVALID, "map_create: bad track length");
        if (track_state[t] > HLSMAP_TRACK_PARSED)
            return fail(HLSMAP_ERR_INVALID, "map_create: bad track_state");
    }
    const int64_t nf = track_offset[n_tracks];
    if (nf > 0 && (!start || !sn)) return fail(HLSMAP_ERR_INVALID, "map_create: null start/sn");

    hlsmap_map *m = new (std::nothrow) hlsmap_map();
    if (!m) return fail(HLSMAP_ERR_NOMEM, "out of host memory");
    m->ctx = ctx;
    m->n_playlists = n_playlists;
    m->n_tracks = n_tracks;
    m->n_frag = nf;
    m->n_frag_padded = ((nf + 3) & ~3ll) + 4;  // bulk copies round spans to 4 elements
    m->h_playlist_base.assign(playlist_track_base, playlist_track_base + n_playlists + 1);
    if (n_playlists > 0 && n_tracks % n_playlists == 0 && (int64_t)n_tracks / n_playlists <= 1024) {
        const int32_t per = n_tracks / n_playlists;
        bool uniform = per > 0;
        for (int32_t p = 0; p < n_playlists && uniform; p++) uniform = playlist_track_base[p] == p * per;
        m->uniform_levels = uniform ? per : 0;
    }
    m->h_url_count.resize(n_tracks);
    for (int32_t t = 0; t < n_tracks; t++) m->h_url_count[t] = url_count ? url_count[t] : 1;

    std::vector<TrackDesc> desc((size_t)n_tracks);
    std::vector<double> dur0((size_t)n_tracks, std::numeric_limits<double>::quiet_NaN());
    int parsed = 0;
    for (int32_t t = 0; t < n_tracks; t++) {
        desc[t].off = track_offset[t];
        desc[t].len = (int)(track_offset[t + 1] - track_offset[t]);
        desc[t].flags = track_state[t];
        if (track_state[t] == HLSMAP_TRACK_PARSED) {
            parsed++;
            if (duration && desc[t].len > 0) dur0[t] = duration[track_offset[t]];
        }
    }

    auto bail = [&](int code) {
        hlsmap_map_destroy(m);
        return code;
    };
#define CKM(call)                                                                             \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            cudaGetLastError();                                                               \
            return bail(fail(e_ == cudaErrorMemoryAllocation ? HLSMAP_ERR_NOMEM : HLSMAP_ERR_CUDA, \
                             std::string(#call) + ": " + cudaGetErrorString(e_)));            \
        }                                                                                     \
    } while (0)

    cudaStream_t s = ctx->stream;
    CKM(cudaMalloc(&m->d_desc, sizeof(TrackDesc) * (size_t)(n_tracks + 1)));
    m->n_tracks_padded = n_tracks > 0 ? n_tracks : 1;
    m->n_base_padded = (n_playlists + 1 + 3) & ~3;
    CKM(cudaMalloc(&m->d_playlist_base, sizeof(int) * (size_t)m->n_base_padded));
    CKM(cudaMemsetAsync(m->d_playlist_base, 0, sizeof(int) * (size_t)m->n_base_padded, ctx->stream));
    CKM(cudaMalloc(&m->d_start, sizeof(double) * (size_t)m->n_frag_padded));
    CKM(cudaMalloc(&m->d_sn, sizeof(int) * (size_t)m->n_frag_padded));
    CKM(cudaMalloc(&m->d_dur0, sizeof(double) * (size_t)(n_tracks + 1)));
    // padding reads as NaN: it can never pass either comparison of the filter
    CKM(cudaMemsetAsync(m->d_start, 0xff, sizeof(double) * (size_t)m->n_frag_padded, s));
    CKM(cudaMemsetAsync(m->d_sn, 0, sizeof(int) * (size_t)m->n_frag_padded, s));
    if (n_tracks) {
        CKM(cudaMemcpyAsync(m->d_desc, desc.data(), sizeof(TrackDesc) * (size_t)n_tracks,
                            cudaMemcpyHostToDevice, s));
        if (dev_dur0)
            CKM(cudaMemcpyAsync(m->d_dur0, dev_dur0, sizeof(double) * (size_t)n_tracks,
                                cudaMemcpyDeviceToDevice, s));
        else
            CKM(cudaMemcpyAsync(m->d_dur0, dur0.data(), sizeof(double) * (size_t)n_tracks,
                                cudaMemcpyHostToDevice, s));
    }
    CKM(cudaMemcpyAsync(m->d_playlist_base, playlist_track_base, sizeof(int) * (size_t)(n_playlists + 1),
                        cudaMemcpyHostToDevice, s));
    if (nf) {
        CKM(cudaMemcpyAsync(m->d_start, start, sizeof(double) * (size_t)nf, cudaMemcpyDefault, s));
        CKM(cudaMemcpyAsync(m->d_sn, sn, sizeof(int) * (size_t)nf, cudaMemcpyDefault, s));
    }
    // monotonicity of every start column, proven on the device (hard part 2)
    CKM(launch_monotone_check(m->d_start, m->d_sn, m->d_desc, n_tracks, s));
    if (n_tracks)
        CKM(cudaMemcpyAsync(desc.data(), m->d_desc, sizeof(TrackDesc) * (size_t)n_tracks,
                            cudaMemcpyDeviceToHost, s));
    CKM(cudaStreamSynchronize(s));

    m->h_monotone.resize(n_tracks);
    m->h_state.assign(track_state, track_state + n_tracks);
    m->h_len.resize(n_tracks);
    for (int32_t t = 0; t < n_tracks; t++) m->h_len[t] = (int32_t)(track_offset[t + 1] - track_offset[t]);
    m->h_cap = m->h_len;
    int mono = 0, contig = 0;
    for (int32_t t = 0; t < n_tracks; t++) {
        m->h_monotone[t] = (desc[t].flags & kTrackMonotone) ? 1 : 0;
        mono += m->h_monotone[t];
        contig += (desc[t].flags & kTrackSnContig) ? 1 : 0;
    }
    m->sn_contig = contig == parsed;
    m->info.n_playlists = n_playlists;
    m->info.n_tracks = n_tracks;
    m->info.n_fragments = nf;
    m->info.n_monotone_tracks = mono;
    m->info.n_parsed_tracks = parsed;
    m->info.all_monotone = (mono == parsed) ? 1 : 0;
    m->info.device_bytes = (int64_t)(sizeof(TrackDesc) * (size_t)(n_tracks + 1) +
                                     sizeof(int) * (size_t)(n_playlists + 1) +
                                     12 * (size_t)m->n_frag_padded + 8 * (size_t)(n_tracks + 1));

    // Shared-memory plan.  SMALL: every start column is sorted and the whole map (start, sn
    // and a u16 hint grid of `cells` cells per fragment) fits next to the kernel's own
    // shared state, so it is staged once per CTA.  Otherwise the search kernel stages, per
    // tile, the window of the start column its queries touch.
    int64_t small_limit = 100 * 1024, window = 64 * 1024, cells = 1;
    if (const char *e = getenv("HLSMAP_SMALL_KB")) small_limit = atoll(e) * 1024;
    if (const char *e = getenv("HLSMAP_STAGE_WINDOW_KB")) window = atoll(e) * 1024;
    if (const char *e = getenv("HLSMAP_HINT_CELLS")) cells = atoll(e) > 0 ? atoll(e) : 1;
    const int64_t smem_budget = (int64_t)ctx->max_smem_optin - 32 * 1024;
    if (small_limit > smem_budget) small_limit = smem_budget;
    if (window > smem_budget) window = smem_budget;
    std::vector<TrackHint> hint;
    int64_t grid_entries = 0;
    bool small = m->info.all_monotone && nf < 65535 && small_limit > 0;
    if (small) {
        hint.resize((size_t)n_tracks);
        for (int32_t t = 0; t < n_tracks; t++) {
            const bool live = (desc[t].flags & kTrackStateMask) == HLSMAP_TRACK_PARSED && desc[t].len > 0;
            hint[t] = TrackHint{};
            hint[t].ncell = live ? (int)(cells * desc[t].len) : 1;
            hint[t].grid_off = (int)grid_entries;
            grid_entries += hint[t].ncell + 1;
        }
        grid_entries = (grid_entries + 7) & ~7ll;
        if (grid_entries < 8) grid_entries = 8;
        if (m->n_frag_padded * 12 + grid_entries * 2 + (int64_t)m->n_tracks_padded * 48 +
                (int64_t)m->n_base_padded * 4 > small_limit)
            small = false;
    }
    if (small) {
        m->grid_padded = (int)grid_entries;
        CKM(cudaMalloc(&m->d_hint, sizeof(TrackHint) * (size_t)(n_tracks + 1)));
        CKM(cudaMalloc(&m->d_grid, sizeof(unsigned short) * (size_t)grid_entries));
        CKM(cudaMemsetAsync(m->d_grid, 0, sizeof(unsigned short) * (size_t)grid_entries, s));
        if (n_tracks)
            CKM(cudaMemcpyAsync(m->d_hint, hint.data(), sizeof(TrackHint) * (size_t)n_tracks,
                                cudaMemcpyHostToDevice, s));
        CKM(launch_build_hint(m->d_start, m->d_desc, m->d_hint, m->d_grid, n_tracks, s));
        CKM(cudaStreamSynchronize(s));
        m->info.stage_all = 1;
        m->stage_elems = (int)m->n_frag_padded;
        m->stage_smem = (size_t)(m->n_frag_padded * 12 + grid_entries * 2 + (int64_t)m->n_tracks_padded * 48 +
                                 (int64_t)m->n_base_padded * 4);
        m->info.device_bytes += (int64_t)(sizeof(TrackHint) * (size_t)(n_tracks + 1) + 2 * (size_t)grid_entries);
    } else {
        m->info.stage_all = 0;
        m->stage_elems = (int)(window / 8) & ~3;
        m->stage_smem = (size_t)m->stage_elems * 8;
    }
    // measured on B200: 512-query tiles win for both kernels (config 2: 4 CTAs/SM instead of 2-3;
    // config 5: a tile's tracks then fit the 64 KB staging window at 3 CTAs/SM, 40 % vs 27 % of
    // the HBM roofline)
    m->ipt = 2;
    if (const char *e = getenv(small ? "HLSMAP_IPT" : "HLSMAP_IPT_GENERAL")) m->ipt = atoi(e) == 4 ? 4 : 2;
    m->grid_search = resolve_search_max_grid(m->stage_smem, small, m->ipt);
    // SMALL maps also offer 1024-query tiles (3 CTAs/SM): picked per batch, see list_core()
    if (small && m->ipt == 2 && !getenv("HLSMAP_IPT")) m->grid_search_wide = resolve_search_max_grid(m->stage_smem, true, 4);
    m->grid_filter = resolve_filter_max_grid();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
        return bail(fail(HLSMAP_ERR_CUDA, std::string("kernel attribute set-up: ") + cudaGetErrorString(e)));
#undef CKM
    *out = m;
    return HLSMAP_OK;
}

int hlsmap_map_create(hlsmap_ctx *ctx, int32_t n_playlists, const int32_t *playlist_track_base,
                      int32_t n_tracks, const int64_t *track_offset, const uint8_t *track_state,
                      const int32_t *url_count, const double *start, const int32_t *sn,
                      const double *duration, hlsmap_map **out)
{
    return map_create_impl(ctx, n_playlists, playlist_track_base, n_tracks, track_offset, track_state,
                           url_count, start, sn, duration, nullptr, out);
}

void hlsmap_map_destroy(hlsmap_map *m)
{
    if (!m) return;
    if (m->ctx) {
        cudaSetDevice(m->ctx->device);
        cudaStreamSynchronize(m->ctx->stream);
    }
    if (m->d_desc) cudaFree(m->d_desc);
    if (m->d_playlist_base) cudaFree(m->d_playlist_base);
    if (m->d_start) cudaFree(m->d_start);
    if (m->d_sn) cudaFree(m->d_sn);
    if (m->d_dur0) cudaFree(m->d_dur0);
    if (m->d_hint) cudaFree(m->d_hint);
    if (m->d_grid) cudaFree(m->d_grid);
    delete m;
}

int hlsmap_map_get_info(const hlsmap_map *map, hlsmap_map_info *info)
{
    if (!map || !info) return fail(HLSMAP_ERR_INVALID, "null argument");
    *info = map->info;
    return HLSMAP_OK;
}

int hlsmap_map_get_monotone(const hlsmap_map *map, uint8_t *flags)
{
    if (!map || !flags) return fail(HLSMAP_ERR_INVALID, "null argument");
    if (map->n_tracks) memcpy(flags, map->h_monotone.data(), (size_t)map->n_tracks);
    return HLSMAP_OK;
}

int hlsmap_map_set_path(hlsmap_map *map, int path)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    if (path != HLSMAP_PATH_AUTO && path != HLSMAP_PATH_SEARCH && path != HLSMAP_PATH_FILTER)
        return fail(HLSMAP_ERR_INVALID, "unknown path");
    if (path == HLSMAP_PATH_SEARCH && !map->info.all_monotone)
        return fail(HLSMAP_ERR_INVALID,
                    "search path refused: a start column is not sorted / NaN-free, only the "
                    "filter kernel is exact for this map");
    map->path = path;
    return HLSMAP_OK;
}

// ---------------------------------------------------------------------------------------
//  live update: slide track windows in place
// ---------------------------------------------------------------------------------------
int hlsmap_map_slide_tracks(hlsmap_map *map, int64_t n, const int32_t *playlist, const int32_t *level,
                            const int32_t *drop, const int64_t *append_offset, const double *new_start,
                            const int32_t *new_sn, const double *new_dur0)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    hlsmap_ctx *ctx = map->ctx;
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && (!level || !drop || !append_offset)))
        return fail(HLSMAP_ERR_INVALID, "slide_tracks: null arrays");
    if (n == 0) return HLSMAP_OK;
    if (!new_dur0) {
        // the duration getSegmentDuration returns is the FIRST fragment's (media-map.js:81-87): evicting
        // at the front, or filling an empty track, changes it, and the reference would read the new one
        for (int64_t i = 0; i < n; i++) {
            const int32_t p = playlist ? playlist[i] : 0;
            const bool known = p >= 0 && p < map->n_playlists && level[i] >= 0 &&
                               level[i] < map->h_playlist_base[p + 1] - map->h_playlist_base[p];
            const bool empty = known && map->h_len[map->h_playlist_base[p] + level[i]] == 0;
            if (drop[i] > 0 || (empty && append_offset[i + 1] > append_offset[i]))
                return fail(HLSMAP_ERR_INVALID, "slide_tracks: new_dur0 is required when a track's first fragment changes "
                                                "(drop > 0, or fragments appended to an empty track)");
        }
    }
    const int64_t n_new = append_offset[n] - append_offset[0];
    if (append_offset[0] != 0 || n_new < 0 || (n_new > 0 && (!new_start || !new_sn)))
        return fail(HLSMAP_ERR_INVALID, "slide_tracks: bad append_offset / null fragment arrays");
    // resolve and validate on the host: a parsed track, at most once, within its reserved size
    std::vector<int32_t> track((size_t)n);
    std::vector<uint8_t> seen((size_t)map->n_tracks, 0);
    for (int64_t i = 0; i < n; i++) {
        const int32_t p = playlist ? playlist[i] : 0;
        if (p < 0 || p >= map->n_playlists) return fail(HLSMAP_ERR_INVALID, "slide_tracks: playlist out of range");
        const int32_t b = map->h_playlist_base[p], e = map->h_playlist_base[p + 1];
        if (level[i] < 0 || level[i] >= e - b) return fail(HLSMAP_ERR_INVALID, "slide_tracks: level out of range");
        const int32_t t = b + level[i];
        if (map->h_state[t] != HLSMAP_TRACK_PARSED) return fail(HLSMAP_ERR_INVALID, "slide_tracks: level is not parsed");
        if (seen[t]) return fail(HLSMAP_ERR_INVALID, "slide_tracks: a track appears twice in one call");
        seen[t] = 1;
        const int64_t add = append_offset[i + 1] - append_offset[i];
        const int32_t k = drop[i] < 0 ? 0 : (drop[i] > map->h_len[t] ? map->h_len[t] : drop[i]);
        if (add < 0 || (int64_t)map->h_len[t] - k + add > map->h_cap[t])
            return fail(HLSMAP_ERR_CAPACITY, "slide_tracks: track would outgrow the " +
                                                 std::to_string(map->h_cap[t]) + " fragments reserved at upload");
        track[i] = t;
    }
    cudaStream_t s = ctx->stream;
    void *v;
    int *d_track, *d_drop, *d_nsn = nullptr;
    long long *d_aoff;
    double *d_nst = nullptr, *d_nd0 = nullptr;
    if ((r = ensure_slot(ctx, 0, n * 4, &v))) return r; d_track = (int *)v;
    if ((r = ensure_slot(ctx, 1, n * 4, &v))) return r; d_drop = (int *)v;
    if ((r = ensure_slot(ctx, 2, (n + 1) * 8, &v))) return r; d_aoff = (long long *)v;
    if ((r = ensure_slot(ctx, 3, (n_new > 0 ? n_new : 1) * 8, &v))) return r; d_nst = (double *)v;
    if ((r = ensure_slot(ctx, 4, (n_new > 0 ? n_new : 1) * 4, &v))) return r; d_nsn = (int *)v;
    if (new_dur0) { if ((r = ensure_slot(ctx, 6, n * 8, &v))) return r; d_nd0 = (double *)v; }
    CK(cudaMemcpyAsync(d_track, track.data(), (size_t)n * 4, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_drop, drop, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_aoff, append_offset, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, s));
    if (n_new > 0) {
        CK(cudaMemcpyAsync(d_nst, new_start, (size_t)n_new * 8, cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(d_nsn, new_sn, (size_t)n_new * 4, cudaMemcpyHostToDevice, s));
    }
    if (new_dur0) CK(cudaMemcpyAsync(d_nd0, new_dur0, (size_t)n * 8, cudaMemcpyHostToDevice, s));
    CK(launch_slide_tracks(map->d_start, map->d_sn, map->d_dur0, map->d_desc, n, d_track, d_drop, d_aoff,
                           d_nst, d_nsn, d_nd0, s));
    // shared-memory-resident maps: the hint grid follows the new columns
    if (map->info.stage_all)
        CK(launch_build_hint(map->d_start, map->d_desc, map->d_hint, map->d_grid, map->n_tracks, s));
    // refresh the host's view of lengths and monotone flags (routes AUTO to search or filter)
    std::vector<TrackDesc> desc((size_t)map->n_tracks);
    CK(cudaMemcpyAsync(desc.data(), map->d_desc, sizeof(TrackDesc) * (size_t)map->n_tracks,
                       cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    int mono = 0, contig = 0;
    int64_t nf = 0;
    for (int32_t t = 0; t < map->n_tracks; t++) {
        map->h_len[t] = desc[t].len;
        map->h_monotone[t] = (desc[t].flags & kTrackMonotone) ? 1 : 0;
        mono += map->h_monotone[t];
        contig += (desc[t].flags & kTrackSnContig) ? 1 : 0;
        nf += desc[t].len;
    }
    map->sn_contig = contig == map->info.n_parsed_tracks;
    map->info.n_fragments = nf;
    map->info.n_monotone_tracks = mono;
    map->info.all_monotone = (mono == map->info.n_parsed_tracks) ? 1 : 0;
    if (!map->info.all_monotone && map->path == HLSMAP_PATH_SEARCH) map->path = HLSMAP_PATH_AUTO;
    return HLSMAP_OK;
}

// ---------------------------------------------------------------------------------------
//  getSegmentList
// ---------------------------------------------------------------------------------------
// One batch on `stream`.  `base` = hits produced by earlier chunks of the same logical call:
// offsets, hit positions and the reported total all include it.
static // The map's tables and the context's scratch, as the list-form kernels take them (after prepare_scan).
void fill_table_params(ResolveParams &p, const hlsmap_map *map, hlsmap_ctx *ctx)
{
    p.desc = map->d_desc;
    p.playlist_base = map->d_playlist_base;
    p.n_playlists = map->n_playlists;
    p.start = map->d_start;
    p.sn = map->d_sn;
    p.n_frag_padded = map->n_frag_padded;
    p.tile_status = status_of(ctx);
    p.ctr = ctr_of(ctx);
    p.result = result_of(ctx);
    p.epoch = ctx->epoch;
    p.tslot = ctx->set_launches[ctx->epoch & 1u];
    p.stage_elems = map->stage_elems;
    p.stage_all = map->info.stage_all;
    p.small = map->info.stage_all;
    p.sn_contig = (map->sn_contig && !map->info.stage_all) ? 1 : 0;
    p.uniform_levels = map->uniform_levels;
    p.hint = map->d_hint;
    p.grid = map->d_grid;
    p.grid_padded = map->grid_padded;
    p.n_tracks_padded = map->n_tracks_padded;
    p.n_base_padded = map->n_base_padded;
}

int list_core(hlsmap_map *map, int64_t nq, const int32_t *d_playlist,
                     const int32_t *d_level, const double *d_begin,
                     const double *d_duration, int32_t *d_offsets, int32_t *d_sn_out,
                     double *d_time_out, int64_t hit_capacity, uint8_t *d_status,
                     void *stream, hlsmap_result *result, unsigned long long base)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    hlsmap_ctx *ctx = map->ctx;
    int r = use_ctx(ctx);
    if (r) return r;
    if (nq < 0 || !d_offsets || (nq > 0 && (!d_level || !d_begin || !d_duration)))
        return fail(HLSMAP_ERR_INVALID, "get_segment_list: null query/offset arrays");
    if ((d_sn_out == nullptr) != (d_time_out == nullptr))
        return fail(HLSMAP_ERR_INVALID, "get_segment_list: sn_out and time_out go together");
    if (hit_capacity < 0) return fail(HLSMAP_ERR_INVALID, "negative hit capacity");
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    const int path = pick_path(map);
    ctx->last_path = path;
    ctx->last_launches = 0;
    ctx->have_last = true;

    if (nq == 0) {
        r = prepare_scan(ctx, 0, s);        // moves to the other scratch set
        if (r) return r;
        mark_begin(ctx, s);
        CK(cudaMemsetAsync(d_offsets, 0, sizeof(int32_t), s));
        CK(cudaMemsetAsync(result_of(ctx), 0, sizeof(DeviceResult), s));
        if (result) return fetch_result(ctx, s, result);
        return HLSMAP_OK;
    }

    // Tile size.  Batches that run back to back (hlsmap_ctx_set_batch_overlap) and give every CTA
    // two tiles or more take 1024-query tiles on SMALL maps: half as many look-backs and head-mark
    // builds per query.  Measured per 10^6-query batch: 35.3 us against 36.3 on the bench mix
    // (12.8 hits per query) and 22.8 against 26.6 on the 30-s look-ahead class (4.6 hits per query),
    // where the per-tile costs weigh most; a batch on its own keeps 512-query tiles (51.7 us
    // against 53.6: its ramp and tail are shorter).
    int ipt = map->ipt;
    if (path == HLSMAP_PATH_SEARCH && ctx->overlap && map->grid_search_wide > 0 &&
        nq >= 2ll * map->grid_search_wide * resolve_search_tile(true, 4))
        ipt = 4;
    const TilePlan plan = plan_tiles(map, path, nq, ipt);
    const int64_t n_tiles = plan.n_tiles;
    if (n_tiles > 0x7fffffff) return fail(HLSMAP_ERR_INVALID, "batch too large: split it");
    r = prepare_scan(ctx, n_tiles, s);
    if (r) return r;

    ResolveParams p{};
    fill_table_params(p, map, ctx);
    p.nq = nq;
    p.q_playlist = d_playlist;
    p.q_level = d_level;
    p.q_begin = d_begin;
    p.q_duration = d_duration;
    p.offsets = d_offsets;
    p.sn_out = d_sn_out;
    p.time_out = d_time_out;
    p.capacity = d_sn_out ? hit_capacity : 0;
    p.status = d_status;
    p.n_tiles = (int)n_tiles;
    p.tile_q = plan.tile_q;
    p.base = base;
    p.ipt = ipt;
    p.vec_ok = al16(d_level) && al16(d_begin) && al16(d_duration) && al16(d_offsets) &&
               (!d_playlist || al16(d_playlist)) &&
               (!d_status || (reinterpret_cast<uintptr_t>(d_status) & 3u) == 0);

    unsigned long long *d_trace = nullptr;
#if HLSMAP_TRACE_BUILD      // diagnostics library only: per-tile phase timestamps into the file HLSMAP_TRACE names
    const char *trace_path = getenv("HLSMAP_TRACE");
    if (trace_path && path == HLSMAP_PATH_SEARCH) {
        CK(cudaMalloc(&d_trace, (size_t)n_tiles * 128));
        CK(cudaMemsetAsync(d_trace, 0, (size_t)n_tiles * 128, s));
        p.trace = d_trace;
    }
#endif
    mark_begin(ctx, s);
    if (path == HLSMAP_PATH_SEARCH) {
        const int max_grid = ipt == map->ipt ? map->grid_search : map->grid_search_wide;
        const int grid = (int)(n_tiles < max_grid ? n_tiles : max_grid);
        CK(launch_resolve_search(p, grid, map->stage_smem, s, ctx->overlap && d_trace == nullptr));
    } else {
        const int grid = (int)(n_tiles < map->grid_filter ? n_tiles : map->grid_filter);
        CK(launch_resolve_filter(p, grid, s));
    }
    ctx->set_launches[ctx->epoch & 1u]++;      // only after a successful launch
    ctx->last_launches = 1;
#if HLSMAP_TRACE_BUILD
    if (d_trace) {
        std::vector<unsigned long long> h((size_t)n_tiles * 16);
        CK(cudaMemcpyAsync(h.data(), d_trace, (size_t)n_tiles * 128, cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        if (FILE *f = fopen(trace_path, "wb")) {
            fwrite(h.data(), 8, h.size(), f);
            fclose(f);
        }
        CK(cudaFree(d_trace));
    }
#endif

    if (result) {
        r = fetch_result(ctx, s, result);
        if (r) return r;
        if (d_sn_out && result->total_hits > hit_capacity)
            return fail(HLSMAP_ERR_CAPACITY, "hit buffers too small: need " +
                                                 std::to_string(result->total_hits) + " entries");
    }
    return HLSMAP_OK;
}

// n >= 2 non-empty batches with ONE launch of the search kernel (their tiles numbered consecutively;
// every CTA stages the tables once for all of them): the batches run concurrently, which the caller
// has allowed with hlsmap_ctx_set_batch_overlap.  Measured on B200, 10^6-query batches on the bench
// map: 33.0 us per batch as overlapped launches, see BASELINE.md for the fused figure.
int list_fused(hlsmap_map *map, const hlsmap_batch *bs, int n, void *stream)
{
    hlsmap_ctx *ctx = map->ctx;
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    int64_t total_q = 0;
    for (int i = 0; i < n; i++) {
        const hlsmap_batch &b = bs[i];
        if (b.nq <= 0 || !b.d_offsets || !b.d_level || !b.d_begin || !b.d_duration)
            return fail(HLSMAP_ERR_INVALID, "get_segment_list: null query/offset arrays");
        if ((b.d_sn_out == nullptr) != (b.d_time_out == nullptr))
            return fail(HLSMAP_ERR_INVALID, "get_segment_list: sn_out and time_out go together");
        if (b.hit_capacity < 0) return fail(HLSMAP_ERR_INVALID, "negative hit capacity");
        total_q += b.nq;
    }
    ctx->last_path = HLSMAP_PATH_SEARCH;
    ctx->last_launches = 0;
    ctx->have_last = true;
    int ipt = map->ipt;
    if (map->grid_search_wide > 0 && total_q >= 2ll * map->grid_search_wide * resolve_search_tile(true, 4)) ipt = 4;
    const int tile_q = resolve_search_tile(map->info.stage_all != 0, ipt);

    ResolveParams p{};
    BatchDesc fb[kMaxFused] = {};
    int64_t n_tiles = 0;
    for (int i = 0; i < n; i++) {
        const hlsmap_batch &b = bs[i];
        BatchDesc &d = fb[i];
        const int64_t t = (b.nq + tile_q - 1) / tile_q;
        d.q_playlist = b.d_playlist;
        d.q_level = b.d_level;
        d.q_begin = b.d_begin;
        d.q_duration = b.d_duration;
        d.offsets = b.d_offsets;
        d.sn_out = b.d_sn_out;
        d.time_out = b.d_time_out;
        d.status = b.d_status;
        d.nq = b.nq;
        d.capacity = b.d_sn_out ? b.hit_capacity : 0;
        d.base = 0ull;
        d.tile_first = (int)n_tiles;
        d.n_tiles = (int)t;
        d.vec_ok = al16(b.d_level) && al16(b.d_begin) && al16(b.d_duration) && al16(b.d_offsets) &&
                   (!b.d_playlist || al16(b.d_playlist)) &&
                   (!b.d_status || (reinterpret_cast<uintptr_t>(b.d_status) & 3u) == 0);
        n_tiles += t;
        if (n_tiles > 0x7fffffff) return fail(HLSMAP_ERR_INVALID, "batches too large: split them");
    }
    int r = prepare_scan(ctx, n_tiles, s);
    if (r) return r;
    fill_table_params(p, map, ctx);
    p.n_tiles = (int)n_tiles;
    p.tile_q = tile_q;
    p.ipt = ipt;
    const int max_grid = ipt == map->ipt ? map->grid_search : map->grid_search_wide;
    const int grid = (int)(n_tiles < max_grid ? n_tiles : max_grid);
    mark_begin(ctx, s);
    CK(launch_resolve_search_fused(p, fb, n, grid, map->stage_smem, s, true));
    ctx->set_launches[ctx->epoch & 1u]++;      // only after a successful launch
    ctx->last_launches = 1;
    return HLSMAP_OK;
}

int hlsmap_get_segment_list_dev(hlsmap_map *map, int64_t nq, const int32_t *d_playlist,
                                const int32_t *d_level, const double *d_begin,
                                const double *d_duration, int32_t *d_offsets, int32_t *d_sn_out,
                                double *d_time_out, int64_t hit_capacity, uint8_t *d_status,
                                void *stream, hlsmap_result *result)
{
    return list_core(map, nq, d_playlist, d_level, d_begin, d_duration, d_offsets, d_sn_out, d_time_out,
                     hit_capacity, d_status, stream, result, 0ull);
}

int hlsmap_get_segment_list_dev_batches(hlsmap_map *map, int32_t n, const hlsmap_batch *batches,
                                        void *stream, int32_t *n_launched)
{
    return hlsmap_get_segment_list_dev_batches_timed(map, n, batches, stream, nullptr, nullptr, n_launched);
}

int hlsmap_get_segment_list_dev_batches_timed(hlsmap_map *map, int32_t n, const hlsmap_batch *batches,
                                              void *stream, void *event_begin, void *event_end,
                                              int32_t *n_launched)
{
    if (n_launched) *n_launched = 0;
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    if (n < 0 || (n > 0 && !batches)) return fail(HLSMAP_ERR_INVALID, "get_segment_list_dev_batches: null batch array");
    {
        int r = use_ctx(map->ctx);
        if (r) return r;
    }
    cudaStream_t s = stream ? (cudaStream_t)stream : map->ctx->stream;
    const bool fuse = map->ctx->overlap && pick_path(map) == HLSMAP_PATH_SEARCH && !map->ctx->no_fuse;
    if (fuse && n > 1) {
        // scratch for the largest launch of the call, before the caller's opening event: growing it
        // means a reallocation, which is set-up and not part of the batches' device time
        const int tile_q = resolve_search_tile(map->info.stage_all != 0, map->ipt);     // (the smaller tile: an upper bound)
        int64_t most = 0, run = 0;
        for (int32_t i = 0, g = 0; i < n; i++) {
            if (batches[i].nq <= 0 || g == kMaxFused) { run = 0; g = 0; }
            if (batches[i].nq > 0) { run += (batches[i].nq + tile_q - 1) / tile_q; g++; }
            if (run > most) most = run;
        }
        if (most <= 0x7fffffff) {
            const int r = reserve_tiles(map->ctx, most, s);
            if (r) return r;
        }
    }
    map->ctx->begin_event = event_begin;     // recorded right before the first launch (mark_begin)
    struct EndMark {        // both events are recorded on every way out
        hlsmap_ctx *ctx;
        void *ev;
        cudaStream_t s;
        ~EndMark()
        {
            mark_begin(ctx, s);               // (no launch happened: an empty or refused call)
            if (ev) cudaEventRecord((cudaEvent_t)ev, s);
        }
    } end_mark{map->ctx, event_end, s};
    // Under hlsmap_ctx_set_batch_overlap the caller has promised that consecutive batches are independent:
    // runs of non-empty batches then go out kMaxFused at a time with one launch each (list_fused).
    int launches = 0;
    for (int32_t i = 0; i < n;) {
        int g = 0;
        if (fuse) {      // (a batch that cannot be launched ends the run before it and fails on its own)
            auto ok = [](const hlsmap_batch &b) {
                return b.nq > 0 && b.d_offsets && b.d_level && b.d_begin && b.d_duration &&
                       (b.d_sn_out == nullptr) == (b.d_time_out == nullptr) && b.hit_capacity >= 0;
            };
            while (i + g < n && g < kMaxFused && ok(batches[i + g])) g++;
        }
        int r;
        if (g >= 2) {
            r = list_fused(map, batches + i, g, stream);
        } else {
            const hlsmap_batch &b = batches[i];
            g = 1;
            r = list_core(map, b.nq, b.d_playlist, b.d_level, b.d_begin, b.d_duration, b.d_offsets,
                          b.d_sn_out, b.d_time_out, b.hit_capacity, b.d_status, stream, nullptr, 0ull);
        }
        if (r) return r;
        launches += map->ctx->last_launches;
        i += g;
        if (n_launched) *n_launched = i;
    }
    if (n > 0) map->ctx->last_launches = launches;     // hlsmap_ctx_last_result: kernel launches of the whole call
    return HLSMAP_OK;
}

int hlsmap_ctx_last_result(hlsmap_ctx *ctx, void *stream, hlsmap_result *result)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (!result) return fail(HLSMAP_ERR_INVALID, "null result");
    if (!ctx->have_last) return fail(HLSMAP_ERR_INVALID, "no batch has been issued");
    return fetch_result(ctx, stream ? (cudaStream_t)stream : ctx->stream, result);
}

int hlsmap_get_segment_list(hlsmap_map *map, int64_t nq, const int32_t *playlist,
                            const int32_t *level, const double *begin, const double *duration,
                            int32_t *offsets, int32_t *sn_out, double *time_out,
                            int64_t hit_capacity, uint8_t *status, hlsmap_result *result)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    hlsmap_ctx *ctx = map->ctx;
    int r = use_ctx(ctx);
    if (r) return r;
    if (nq < 0 || !offsets || (nq > 0 && (!level || !begin || !duration)))
        return fail(HLSMAP_ERR_INVALID, "get_segment_list: null query/offset arrays");
    if ((sn_out == nullptr) != (time_out == nullptr))
        return fail(HLSMAP_ERR_INVALID, "get_segment_list: sn_out and time_out go together");
    cudaStream_t s = ctx->stream;
    const bool fill = sn_out != nullptr;

    int *d_pl = nullptr, *d_lv = nullptr, *d_off = nullptr, *d_sn = nullptr;
    double *d_bg = nullptr, *d_du = nullptr, *d_tm = nullptr;
    uint8_t *d_st = nullptr;
    void *v;
    const int64_t n1 = nq > 0 ? nq : 1;
    if (playlist) { if ((r = ensure_slot(ctx, 0, n1 * 4, &v))) return r; d_pl = (int *)v; }
    if ((r = ensure_slot(ctx, 1, n1 * 4, &v))) return r; d_lv = (int *)v;
    if ((r = ensure_slot(ctx, 2, n1 * 8, &v))) return r; d_bg = (double *)v;
    if ((r = ensure_slot(ctx, 3, n1 * 8, &v))) return r; d_du = (double *)v;
    if ((r = ensure_slot(ctx, 4, (n1 + 1) * 4, &v))) return r; d_off = (int *)v;
    if (status) { if ((r = ensure_slot(ctx, 5, n1, &v))) return r; d_st = (uint8_t *)v; }
    // Large filling batches go through a chunked pipeline: while chunk c's hits travel to the
    // host, chunk c+1's queries travel to the device (PCIe is full duplex) and its kernel runs.
    // Chunks of host_chunk queries (a multiple of 1024: aligned sections).  Few, large copies win:
    // measured per 10^6-query step 3.10 ms with 262144-query chunks, 3.19 with 131072, 3.39 with
    // 65536, 3.29 without the pipeline; sizes ramping up from 32768 (the first hits on their way
    // back sooner, but two more chunks) 3.28.
    const int64_t kChunk = ctx->host_chunk;
    if (fill && nq >= 2 * kChunk && ctx->host_pipeline) {
        std::vector<int64_t> cut{0};
        while (cut.back() < nq) cut.push_back(cut.back() + kChunk < nq ? cut.back() + kChunk : nq);
        int64_t dev_cap = ctx->slot_bytes[6] / 4 < ctx->slot_bytes[7] / 8 ? ctx->slot_bytes[6] / 4
                                                                         : ctx->slot_bytes[7] / 8;
        int64_t want = nq * 16 + 1024;
        if (want > hit_capacity) want = hit_capacity;
        if (dev_cap < want) dev_cap = want;
        if (dev_cap > hit_capacity) dev_cap = hit_capacity;
        if ((r = ensure_slot(ctx, 6, dev_cap * 4, &v))) return r; d_sn = (int *)v;
        if ((r = ensure_slot(ctx, 7, dev_cap * 8, &v))) return r; d_tm = (double *)v;
        auto upload = [&](int64_t c) -> int {
            const int64_t q0 = cut[c], n = cut[c + 1] - cut[c];
            if (playlist) CK(cudaMemcpyAsync(d_pl + q0, playlist + q0, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->s_in));
            CK(cudaMemcpyAsync(d_lv + q0, level + q0, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->s_in));
            CK(cudaMemcpyAsync(d_bg + q0, begin + q0, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->s_in));
            CK(cudaMemcpyAsync(d_du + q0, duration + q0, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->s_in));
            CK(cudaEventRecord(ctx->ev_in[c & 1], ctx->s_in));
            return HLSMAP_OK;
        };
        const int64_t n_chunks = (int64_t)cut.size() - 1;
        hlsmap_result acc{};
        unsigned long long base = 0;
        bool overflow_dev = false;
        CK(cudaStreamSynchronize(ctx->s_out));          // earlier downloads are done with the slots
        if ((r = upload(0))) return r;
        for (int64_t c = 0; c < n_chunks; c++) {
            const int64_t q0 = cut[c], n = cut[c + 1] - cut[c];
            CK(cudaStreamWaitEvent(s, ctx->ev_in[c & 1], 0));
            if (c + 1 < n_chunks && (r = upload(c + 1))) return r;
            hlsmap_result rc{};
            r = list_core(map, n, playlist ? d_pl + q0 : nullptr, d_lv + q0, d_bg + q0, d_du + q0,
                          d_off + q0, d_sn, d_tm, dev_cap, status ? d_st + q0 : nullptr, s, &rc, base);
            acc.kernel_launches += 1;
            acc.path = rc.path;
            if (r == HLSMAP_ERR_CAPACITY) { overflow_dev = true; break; }   // device guess too small
            if (r) return r;
            const unsigned long long total_c = (unsigned long long)rc.total_hits - base;
            acc.n_unparsed += rc.n_unparsed;
            acc.n_missing += rc.n_missing;
            CK(cudaMemcpyAsync(offsets + q0, d_off + q0, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->s_out));
            if (status) CK(cudaMemcpyAsync(status + q0, d_st + q0, (size_t)n, cudaMemcpyDeviceToHost, ctx->s_out));
            if (total_c > 0 && (int64_t)rc.total_hits <= hit_capacity) {
                CK(cudaMemcpyAsync(sn_out + base, d_sn + base, (size_t)total_c * 4, cudaMemcpyDeviceToHost, ctx->s_out));
                CK(cudaMemcpyAsync(time_out + base, d_tm + base, (size_t)total_c * 8, cudaMemcpyDeviceToHost, ctx->s_out));
            }
            base = (unsigned long long)rc.total_hits;
        }
        CK(cudaStreamSynchronize(ctx->s_in));
        CK(cudaStreamSynchronize(ctx->s_out));
        if (!overflow_dev) {
            offsets[nq] = (int32_t)base;
            acc.total_hits = (int64_t)base;
            ctx->last_launches = acc.kernel_launches;
            if (result) *result = acc;
            if ((int64_t)base > hit_capacity)
                return fail(HLSMAP_ERR_CAPACITY, "hit buffers too small: need " + std::to_string(base) + " entries");
            return HLSMAP_OK;
        }
        // the device-side guess was too small: fall through to the one-shot path, which sizes
        // the device buffers from the exact count
    }
    if (nq) {
        if (playlist) CK(cudaMemcpyAsync(d_pl, playlist, (size_t)nq * 4, cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(d_lv, level, (size_t)nq * 4, cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(d_bg, begin, (size_t)nq * 8, cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(d_du, duration, (size_t)nq * 8, cudaMemcpyHostToDevice, s));
    }

    hlsmap_result res{};
    int launches = 0;
    int64_t dev_cap = 0;
    if (fill) {
        // device hit buffers: grow-only; first guess 8 hits per query, bounded by the caller
        dev_cap = ctx->slot_bytes[6] / 4 < ctx->slot_bytes[7] / 8 ? ctx->slot_bytes[6] / 4
                                                                 : ctx->slot_bytes[7] / 8;
        int64_t want = nq * 8 + 1024;
        if (want > hit_capacity) want = hit_capacity;
        if (dev_cap < want) dev_cap = want;
        if (dev_cap > hit_capacity) dev_cap = hit_capacity;
        if ((r = ensure_slot(ctx, 6, dev_cap * 4, &v))) return r; d_sn = (int *)v;
        if ((r = ensure_slot(ctx, 7, dev_cap * 8, &v))) return r; d_tm = (double *)v;
    }
    r = hlsmap_get_segment_list_dev(map, nq, d_pl, d_lv, d_bg, d_du, d_off, d_sn, d_tm, dev_cap,
                                    d_st, s, &res);
    launches += ctx->last_launches;
    if (r == HLSMAP_ERR_CAPACITY && res.total_hits <= hit_capacity) {
        // the guess was too small but the caller's buffers are not: grow and rerun
        dev_cap = res.total_hits;
        if ((r = ensure_slot(ctx, 6, dev_cap * 4, &v))) return r; d_sn = (int *)v;
        if ((r = ensure_slot(ctx, 7, dev_cap * 8, &v))) return r; d_tm = (double *)v;
        r = hlsmap_get_segment_list_dev(map, nq, d_pl, d_lv, d_bg, d_du, d_off, d_sn, d_tm,
                                        dev_cap, d_st, s, &res);
        launches += ctx->last_launches;
    }
    res.kernel_launches = launches;
    ctx->last_launches = launches;
    if (result) *result = res;
    if (r != HLSMAP_OK && r != HLSMAP_ERR_CAPACITY) return r;

    CK(cudaMemcpyAsync(offsets, d_off, (size_t)(nq + 1) * 4, cudaMemcpyDeviceToHost, s));
    if (status && nq) CK(cudaMemcpyAsync(status, d_st, (size_t)nq, cudaMemcpyDeviceToHost, s));
    const bool fits = fill && res.total_hits <= hit_capacity;
    if (fits && res.total_hits > 0) {
        CK(cudaMemcpyAsync(sn_out, d_sn, (size_t)res.total_hits * 4, cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(time_out, d_tm, (size_t)res.total_hits * 8, cudaMemcpyDeviceToHost, s));
    }
    CK(cudaStreamSynchronize(s));
    if (fill && !fits)
        return fail(HLSMAP_ERR_CAPACITY, "hit buffers too small: need " +
                                             std::to_string(res.total_hits) + " entries");
    return HLSMAP_OK;
}

// ---------------------------------------------------------------------------------------
//  range form
// ---------------------------------------------------------------------------------------
int hlsmap_get_segment_range_dev(hlsmap_map *map, int64_t nq, const int32_t *d_playlist,
                                 const int32_t *d_level, const double *d_begin,
                                 const double *d_duration, int32_t *d_first, int32_t *d_count,
                                 uint8_t *d_status, void *stream)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    hlsmap_ctx *ctx = map->ctx;
    int r = use_ctx(ctx);
    if (r) return r;
    if (nq < 0 || (nq > 0 && (!d_level || !d_begin || !d_duration || !d_first || !d_count)))
        return fail(HLSMAP_ERR_INVALID, "get_segment_range: null arrays");
    if (nq == 0) return HLSMAP_OK;
    RangeParams p{};
    p.desc = map->d_desc;
    p.playlist_base = map->d_playlist_base;
    p.n_playlists = map->n_playlists;
    p.start = map->d_start;
    p.nq = nq;
    p.q_playlist = d_playlist;
    p.q_level = d_level;
    p.q_begin = d_begin;
    p.q_duration = d_duration;
    p.first = d_first;
    p.count = d_count;
    p.status = d_status;
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    if (map->info.stage_all && pick_path(map) == HLSMAP_PATH_SEARCH) {
        // the whole map fits shared memory and every column is sorted: staged tables + hint grid
        const int vec_ok = al16(d_level) && al16(d_begin) && al16(d_duration) && al16(d_first) && al16(d_count) &&
                           (!d_playlist || al16(d_playlist)) &&
                           (!d_status || (reinterpret_cast<uintptr_t>(d_status) & 3u) == 0);
        CK(launch_range_small(p, map->d_hint, map->d_grid, map->n_frag_padded, map->grid_padded,
                              map->n_tracks_padded, map->n_base_padded, vec_ok, s));
    } else {
        CK(launch_range(p, s));
    }
    ctx->last_launches = 1;
    return HLSMAP_OK;
}

int hlsmap_get_segment_range(hlsmap_map *map, int64_t nq, const int32_t *playlist,
                             const int32_t *level, const double *begin, const double *duration,
                             int32_t *first, int32_t *count, uint8_t *status)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    hlsmap_ctx *ctx = map->ctx;
    int r = use_ctx(ctx);
    if (r) return r;
    if (nq < 0 || (nq > 0 && (!level || !begin || !duration || !first || !count)))
        return fail(HLSMAP_ERR_INVALID, "get_segment_range: null arrays");
    if (nq == 0) return HLSMAP_OK;
    cudaStream_t s = ctx->stream;
    void *v;
    int *d_pl = nullptr, *d_lv, *d_first, *d_count;
    double *d_bg, *d_du;
    uint8_t *d_st = nullptr;
    if (playlist) { if ((r = ensure_slot(ctx, 0, nq * 4, &v))) return r; d_pl = (int *)v; }
    if ((r = ensure_slot(ctx, 1, nq * 4, &v))) return r; d_lv = (int *)v;
    if ((r = ensure_slot(ctx, 2, nq * 8, &v))) return r; d_bg = (double *)v;
    if ((r = ensure_slot(ctx, 3, nq * 8, &v))) return r; d_du = (double *)v;
    if ((r = ensure_slot(ctx, 4, nq * 4, &v))) return r; d_first = (int *)v;
    if ((r = ensure_slot(ctx, 6, nq * 4, &v))) return r; d_count = (int *)v;
    if (status) { if ((r = ensure_slot(ctx, 5, nq, &v))) return r; d_st = (uint8_t *)v; }
    if (playlist) CK(cudaMemcpyAsync(d_pl, playlist, (size_t)nq * 4, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_lv, level, (size_t)nq * 4, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_bg, begin, (size_t)nq * 8, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_du, duration, (size_t)nq * 8, cudaMemcpyHostToDevice, s));
    r = hlsmap_get_segment_range_dev(map, nq, d_pl, d_lv, d_bg, d_du, d_first, d_count, d_st, s);
    if (r) return r;
    CK(cudaMemcpyAsync(first, d_first, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(count, d_count, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
    if (status) CK(cudaMemcpyAsync(status, d_st, (size_t)nq, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return HLSMAP_OK;
}

// ---------------------------------------------------------------------------------------
//  getSegmentDuration / getTrackList
// ---------------------------------------------------------------------------------------
int hlsmap_get_segment_duration_dev(hlsmap_map *map, int64_t n, const int32_t *d_playlist,
                                    const int32_t *d_level, double *d_out, uint8_t *d_status,
                                    void *stream)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    hlsmap_ctx *ctx = map->ctx;
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && (!d_level || !d_out)))
        return fail(HLSMAP_ERR_INVALID, "get_segment_duration: null arrays");
    if (n == 0) return HLSMAP_OK;
    CK(launch_segment_duration(map->d_desc, map->d_playlist_base, map->n_playlists, map->d_dur0, n,
                               d_playlist, d_level, d_out, d_status,
                               stream ? (cudaStream_t)stream : ctx->stream));
    ctx->last_launches = 1;
    return HLSMAP_OK;
}

int hlsmap_get_segment_duration(hlsmap_map *map, int64_t n, const int32_t *playlist,
                                const int32_t *level, double *out, uint8_t *status)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    hlsmap_ctx *ctx = map->ctx;
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && (!level || !out)))
        return fail(HLSMAP_ERR_INVALID, "get_segment_duration: null arrays");
    if (n == 0) return HLSMAP_OK;
    cudaStream_t s = ctx->stream;
    void *v;
    int *d_pl = nullptr, *d_lv;
    double *d_out;
    uint8_t *d_st = nullptr;
    if (playlist) { if ((r = ensure_slot(ctx, 0, n * 4, &v))) return r; d_pl = (int *)v; }
    if ((r = ensure_slot(ctx, 1, n * 4, &v))) return r; d_lv = (int *)v;
    if ((r = ensure_slot(ctx, 2, n * 8, &v))) return r; d_out = (double *)v;
    if (status) { if ((r = ensure_slot(ctx, 5, n, &v))) return r; d_st = (uint8_t *)v; }
    if (playlist) CK(cudaMemcpyAsync(d_pl, playlist, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_lv, level, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    r = hlsmap_get_segment_duration_dev(map, n, d_pl, d_lv, d_out, d_st, s);
    if (r) return r;
    CK(cudaMemcpyAsync(out, d_out, (size_t)n * 8, cudaMemcpyDeviceToHost, s));
    if (status) CK(cudaMemcpyAsync(status, d_st, (size_t)n, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return HLSMAP_OK;
}

// ---- views from fragments / next segment view (extension) -----------------------------------
int hlsmap_views_from_fragments_dev(hlsmap_map *map, int64_t n, const int32_t *d_playlist,
                                    const int32_t *d_level, const int32_t *d_index, int32_t *d_sn_out,
                                    double *d_time_out, uint8_t *d_status, void *stream)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    hlsmap_ctx *ctx = map->ctx;
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && (!d_level || !d_index || !d_sn_out || !d_time_out)))
        return fail(HLSMAP_ERR_INVALID, "views_from_fragments: null arrays");
    if (n == 0) return HLSMAP_OK;
    CK(launch_views_from_fragments(map->d_desc, map->d_playlist_base, map->n_playlists, map->d_start,
                                   map->d_sn, n, d_playlist, d_level, d_index, d_sn_out, d_time_out,
                                   d_status, stream ? (cudaStream_t)stream : ctx->stream));
    ctx->last_launches = 1;
    return HLSMAP_OK;
}

int hlsmap_get_next_segment_view_dev(hlsmap_map *map, int64_t n, const int32_t *d_playlist,
                                     const int32_t *d_level, const int32_t *d_sn, int32_t *d_sn_out,
                                     double *d_time_out, uint8_t *d_found, uint8_t *d_status,
                                     void *stream)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    hlsmap_ctx *ctx = map->ctx;
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && (!d_level || !d_sn || !d_sn_out || !d_time_out || !d_found)))
        return fail(HLSMAP_ERR_INVALID, "get_next_segment_view: null arrays");
    if (n == 0) return HLSMAP_OK;
    CK(launch_next_segment_view(map->d_desc, map->d_playlist_base, map->n_playlists, map->d_start,
                                map->d_sn, n, d_playlist, d_level, d_sn, d_sn_out, d_time_out, d_found,
                                d_status, stream ? (cudaStream_t)stream : ctx->stream));
    ctx->last_launches = 1;
    return HLSMAP_OK;
}

// shared host-pointer wrapper: 3 int32 inputs, (sn, time[, found], status) outputs
static int frag_host_call(hlsmap_map *map, int64_t n, const int32_t *playlist, const int32_t *level,
                          const int32_t *third, int32_t *sn_out, double *time_out, uint8_t *found,
                          uint8_t *status, bool next_view)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    hlsmap_ctx *ctx = map->ctx;
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && (!level || !third || !sn_out || !time_out || (next_view && !found))))
        return fail(HLSMAP_ERR_INVALID, "null arrays");
    if (n == 0) return HLSMAP_OK;
    cudaStream_t s = ctx->stream;
    void *v;
    int *d_pl = nullptr, *d_lv, *d_th, *d_sn;
    double *d_tm;
    uint8_t *d_st = nullptr, *d_fd = nullptr;
    if (playlist) { if ((r = ensure_slot(ctx, 0, n * 4, &v))) return r; d_pl = (int *)v; }
    if ((r = ensure_slot(ctx, 1, n * 4, &v))) return r; d_lv = (int *)v;
    if ((r = ensure_slot(ctx, 4, n * 4, &v))) return r; d_th = (int *)v;
    if ((r = ensure_slot(ctx, 6, n * 4, &v))) return r; d_sn = (int *)v;
    if ((r = ensure_slot(ctx, 7, n * 8, &v))) return r; d_tm = (double *)v;
    if (status) { if ((r = ensure_slot(ctx, 5, n, &v))) return r; d_st = (uint8_t *)v; }
    if (next_view) { if ((r = ensure_slot(ctx, 8, n, &v))) return r; d_fd = (uint8_t *)v; }
    if (playlist) CK(cudaMemcpyAsync(d_pl, playlist, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_lv, level, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_th, third, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    r = next_view ? hlsmap_get_next_segment_view_dev(map, n, d_pl, d_lv, d_th, d_sn, d_tm, d_fd, d_st, s)
                  : hlsmap_views_from_fragments_dev(map, n, d_pl, d_lv, d_th, d_sn, d_tm, d_st, s);
    if (r) return r;
    CK(cudaMemcpyAsync(sn_out, d_sn, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(time_out, d_tm, (size_t)n * 8, cudaMemcpyDeviceToHost, s));
    if (status) CK(cudaMemcpyAsync(status, d_st, (size_t)n, cudaMemcpyDeviceToHost, s));
    if (next_view) CK(cudaMemcpyAsync(found, d_fd, (size_t)n, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return HLSMAP_OK;
}

int hlsmap_views_from_fragments(hlsmap_map *map, int64_t n, const int32_t *playlist,
                                const int32_t *level, const int32_t *index, int32_t *sn_out,
                                double *time_out, uint8_t *status)
{
    return frag_host_call(map, n, playlist, level, index, sn_out, time_out, nullptr, status, false);
}

int hlsmap_get_next_segment_view(hlsmap_map *map, int64_t n, const int32_t *playlist,
                                 const int32_t *level, const int32_t *sn, int32_t *sn_out,
                                 double *time_out, uint8_t *found, uint8_t *status)
{
    return frag_host_call(map, n, playlist, level, sn, sn_out, time_out, found, status, true);
}

int hlsmap_get_track_list(const hlsmap_map *map, int32_t playlist, int32_t *level_out,
                          int32_t *url_id_out, int64_t capacity, int64_t *n_out)
{
    if (!map || !n_out) return fail(HLSMAP_ERR_INVALID, "null argument");
    if (playlist < 0 || playlist >= map->n_playlists)
        return fail(HLSMAP_ERR_INVALID, "playlist out of range");
    if ((level_out == nullptr) != (url_id_out == nullptr))
        return fail(HLSMAP_ERR_INVALID, "level_out and url_id_out go together");
    const int32_t b = map->h_playlist_base[playlist], e = map->h_playlist_base[playlist + 1];
    int64_t k = 0;
    for (int32_t t = b; t < e; t++)              // for i in levels          (media-map.js:66)
        for (int32_t j = 0; j < map->h_url_count[t]; j++) {  // for j in level.url (:68)
            if (level_out && k < capacity) {
                level_out[k] = t - b;
                url_id_out[k] = j;
            }
            k++;
        }
    *n_out = k;
    if (level_out && k > capacity) return fail(HLSMAP_ERR_CAPACITY, "track list buffers too small");
    return HLSMAP_OK;
}

// ---------------------------------------------------------------------------------------
//  views
// ---------------------------------------------------------------------------------------
#define VIEW_PRELUDE(cond)                                                     \
    int r = use_ctx(ctx);                                                      \
    if (r) return r;                                                           \
    if (n < 0 || (n > 0 && !(cond))) return fail(HLSMAP_ERR_INVALID, "null view arrays"); \
    if (n == 0) return HLSMAP_OK;                                              \
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;              \
    ctx->last_launches = 1;

int hlsmap_view_pack_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *d_level,
                         const int32_t *d_url_id, const int32_t *d_sn, uint32_t *d_wire,
                         void *stream)
{
    VIEW_PRELUDE(d_level && d_url_id && d_sn && d_wire)
    CK(launch_view_pack(n, d_level, d_url_id, d_sn, d_wire, s));
    return HLSMAP_OK;
}

int hlsmap_view_unpack_dev(hlsmap_ctx *ctx, int64_t n, const uint32_t *d_wire, uint32_t *d_level,
                           uint32_t *d_url_id, uint32_t *d_sn, void *stream)
{
    VIEW_PRELUDE(d_level && d_url_id && d_sn && d_wire)
    CK(launch_view_unpack(n, d_wire, d_level, d_url_id, d_sn, s));
    return HLSMAP_OK;
}

int hlsmap_view_is_equal_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *al, const int32_t *au,
                             const int32_t *as, const int32_t *bl, const int32_t *bu,
                             const int32_t *bs, uint8_t *d_out, void *stream)
{
    VIEW_PRELUDE(al && au && as && bl && bu && bs && d_out)
    CK(launch_view_is_equal(n, al, au, as, bl, bu, bs, d_out, s));
    return HLSMAP_OK;
}

int hlsmap_view_is_in_track_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *d_level,
                                const int32_t *d_url_id, const int32_t *d_t_level,
                                const int32_t *d_t_url_id, uint8_t *d_out, void *stream)
{
    VIEW_PRELUDE(d_level && d_url_id && d_t_level && d_t_url_id && d_out)
    CK(launch_view_is_in_track(n, d_level, d_url_id, d_t_level, d_t_url_id, d_out, s));
    return HLSMAP_OK;
}

int hlsmap_view_roundtrip_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *d_level,
                              const int32_t *d_url_id, const int32_t *d_sn, uint32_t *d_wire,
                              uint8_t *d_equal, void *stream)
{
    VIEW_PRELUDE(d_level && d_url_id && d_sn && d_wire && d_equal)
    CK(launch_view_roundtrip(n, d_level, d_url_id, d_sn, d_wire, d_equal, s));
    return HLSMAP_OK;
}

int hlsmap_view_format_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *d_level,
                           const int32_t *d_url_id, const int32_t *d_sn, int with_sn,
                           int32_t *d_offsets, char *d_chars, int64_t capacity, void *stream,
                           int64_t *total_out)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || !d_offsets || capacity < 0 ||
        (n > 0 && (!d_level || !d_url_id || (with_sn && !d_sn) || (capacity > 0 && !d_chars))))
        return fail(HLSMAP_ERR_INVALID, "view_format: null arrays");
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    ctx->have_last = true;
    ctx->last_path = 0;
    ctx->last_launches = 0;
    if (n == 0) {
        CK(cudaMemsetAsync(d_offsets, 0, sizeof(int32_t), s));
        if (total_out) *total_out = 0;
        return HLSMAP_OK;
    }
    const int64_t n_tiles = (n + kFormatTileViews - 1) / kFormatTileViews;
    if (n_tiles > 0x7fffffff) return fail(HLSMAP_ERR_INVALID, "batch too large: split it");
    static const bool single_pass = getenv("HLSMAP_FORMAT_SINGLE_PASS") != nullptr;   // diagnostics: the look-back form
    void *st_b = nullptr;
    if (single_pass) r = prepare_scan(ctx, n_tiles, s);
    else r = ensure_slot(ctx, 10, (n_tiles + 1) * 16, &st_b, s);
    if (r) return r;
    FormatParams p{};
    p.n = n;
    p.level = d_level;
    p.url_id = d_url_id;
    p.sn = d_sn;
    p.with_sn = with_sn ? 1 : 0;
    p.offsets = d_offsets;
    p.chars = d_chars;
    p.capacity = capacity;
    p.tile_status = status_of(ctx);
    p.ctr = ctr_of(ctx);
    p.result = result_of(ctx);
    p.epoch = ctx->epoch;
    p.tslot = ctx->set_launches[ctx->epoch & 1u];
    p.n_tiles = (int)n_tiles;
    if (single_pass) {
        const int maxg = view_format_max_grid();
        CK(launch_view_format(p, (int)(n_tiles < maxg ? n_tiles : maxg), s));
        ctx->set_launches[ctx->epoch & 1u]++;
        ctx->last_launches = 1;
    } else {
        p.block_sum = (unsigned long long *)st_b;
        p.block_base = (unsigned long long *)st_b + n_tiles + 1;
        CK(launch_view_format_split(p, s));
        ctx->last_launches = 3;
    }
    if (total_out) {
        hlsmap_result res{};
        r = fetch_result(ctx, s, &res);
        if (r) return r;
        *total_out = res.total_hits;
        if (res.total_hits > capacity)
            return fail(HLSMAP_ERR_CAPACITY, "character buffer too small: need " +
                                                 std::to_string(res.total_hits) + " bytes");
    }
    return HLSMAP_OK;
}

// ---- host-pointer view variants: copy in, run the kernel, copy out -----------------------
namespace {
struct Cols {
    const void *h;
    int64_t bytes;
    void *d;
};
int stage_in(hlsmap_ctx *ctx, Cols *c, int n, int first_slot)
{
    for (int i = 0; i < n; i++) {
        int r = ensure_slot(ctx, first_slot + i, c[i].bytes, &c[i].d);
        if (r) return r;
        if (c[i].h) CK(cudaMemcpyAsync(c[i].d, c[i].h, (size_t)c[i].bytes, cudaMemcpyHostToDevice, ctx->stream));
    }
    return HLSMAP_OK;
}
}  // namespace

int hlsmap_view_pack(hlsmap_ctx *ctx, int64_t n, const int32_t *level, const int32_t *url_id,
                     const int32_t *sn, uint32_t *wire)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && !(level && url_id && sn && wire)))
        return fail(HLSMAP_ERR_INVALID, "null view arrays");
    if (n == 0) return HLSMAP_OK;
    Cols c[4] = {{level, n * 4, nullptr}, {url_id, n * 4, nullptr}, {sn, n * 4, nullptr},
                 {nullptr, n * 12, nullptr}};
    if ((r = stage_in(ctx, c, 4, 0))) return r;
    r = hlsmap_view_pack_dev(ctx, n, (int *)c[0].d, (int *)c[1].d, (int *)c[2].d, (uint32_t *)c[3].d,
                             ctx->stream);
    if (r) return r;
    CK(cudaMemcpyAsync(wire, c[3].d, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return HLSMAP_OK;
}

int hlsmap_view_unpack(hlsmap_ctx *ctx, int64_t n, const uint32_t *wire, uint32_t *level,
                       uint32_t *url_id, uint32_t *sn)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && !(level && url_id && sn && wire)))
        return fail(HLSMAP_ERR_INVALID, "null view arrays");
    if (n == 0) return HLSMAP_OK;
    Cols c[4] = {{wire, n * 12, nullptr}, {nullptr, n * 4, nullptr}, {nullptr, n * 4, nullptr},
                 {nullptr, n * 4, nullptr}};
    if ((r = stage_in(ctx, c, 4, 0))) return r;
    r = hlsmap_view_unpack_dev(ctx, n, (uint32_t *)c[0].d, (uint32_t *)c[1].d, (uint32_t *)c[2].d,
                               (uint32_t *)c[3].d, ctx->stream);
    if (r) return r;
    CK(cudaMemcpyAsync(level, c[1].d, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(url_id, c[2].d, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(sn, c[3].d, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return HLSMAP_OK;
}

int hlsmap_view_is_equal(hlsmap_ctx *ctx, int64_t n, const int32_t *a_level,
                         const int32_t *a_url_id, const int32_t *a_sn, const int32_t *b_level,
                         const int32_t *b_url_id, const int32_t *b_sn, uint8_t *out)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && !(a_level && a_url_id && a_sn && b_level && b_url_id && b_sn && out)))
        return fail(HLSMAP_ERR_INVALID, "null view arrays");
    if (n == 0) return HLSMAP_OK;
    Cols c[7] = {{a_level, n * 4, nullptr}, {a_url_id, n * 4, nullptr}, {a_sn, n * 4, nullptr},
                 {b_level, n * 4, nullptr}, {b_url_id, n * 4, nullptr}, {b_sn, n * 4, nullptr},
                 {nullptr, n, nullptr}};
    if ((r = stage_in(ctx, c, 7, 0))) return r;
    r = hlsmap_view_is_equal_dev(ctx, n, (int *)c[0].d, (int *)c[1].d, (int *)c[2].d, (int *)c[3].d,
                                 (int *)c[4].d, (int *)c[5].d, (uint8_t *)c[6].d, ctx->stream);
    if (r) return r;
    CK(cudaMemcpyAsync(out, c[6].d, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return HLSMAP_OK;
}

int hlsmap_view_is_in_track(hlsmap_ctx *ctx, int64_t n, const int32_t *level,
                            const int32_t *url_id, const int32_t *t_level,
                            const int32_t *t_url_id, uint8_t *out)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && !(level && url_id && t_level && t_url_id && out)))
        return fail(HLSMAP_ERR_INVALID, "null view arrays");
    if (n == 0) return HLSMAP_OK;
    Cols c[5] = {{level, n * 4, nullptr}, {url_id, n * 4, nullptr}, {t_level, n * 4, nullptr},
                 {t_url_id, n * 4, nullptr}, {nullptr, n, nullptr}};
    if ((r = stage_in(ctx, c, 5, 0))) return r;
    r = hlsmap_view_is_in_track_dev(ctx, n, (int *)c[0].d, (int *)c[1].d, (int *)c[2].d,
                                    (int *)c[3].d, (uint8_t *)c[4].d, ctx->stream);
    if (r) return r;
    CK(cudaMemcpyAsync(out, c[4].d, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return HLSMAP_OK;
}

int hlsmap_view_roundtrip(hlsmap_ctx *ctx, int64_t n, const int32_t *level,
                          const int32_t *url_id, const int32_t *sn, uint32_t *wire,
                          uint8_t *equal)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && !(level && url_id && sn && wire && equal)))
        return fail(HLSMAP_ERR_INVALID, "null view arrays");
    if (n == 0) return HLSMAP_OK;
    Cols c[5] = {{level, n * 4, nullptr}, {url_id, n * 4, nullptr}, {sn, n * 4, nullptr},
                 {nullptr, n * 12, nullptr}, {nullptr, n, nullptr}};
    if ((r = stage_in(ctx, c, 5, 0))) return r;
    r = hlsmap_view_roundtrip_dev(ctx, n, (int *)c[0].d, (int *)c[1].d, (int *)c[2].d,
                                  (uint32_t *)c[3].d, (uint8_t *)c[4].d, ctx->stream);
    if (r) return r;
    CK(cudaMemcpyAsync(wire, c[3].d, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(equal, c[4].d, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return HLSMAP_OK;
}

int hlsmap_view_format(hlsmap_ctx *ctx, int64_t n, const int32_t *level, const int32_t *url_id,
                       const int32_t *sn, int with_sn, int32_t *offsets, char *chars,
                       int64_t capacity, int64_t *total_out)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || !offsets || capacity < 0 ||
        (n > 0 && (!level || !url_id || (with_sn && !sn) || (capacity > 0 && !chars))))
        return fail(HLSMAP_ERR_INVALID, "view_format: null arrays");
    const int64_t n1 = n > 0 ? n : 1;
    Cols c[5] = {{level, n1 * 4, nullptr}, {url_id, n1 * 4, nullptr},
                 {with_sn ? sn : nullptr, n1 * 4, nullptr}, {nullptr, (n1 + 1) * 4, nullptr},
                 {nullptr, capacity > 0 ? capacity : 16, nullptr}};
    if (n == 0) c[0].h = c[1].h = c[2].h = nullptr;
    if ((r = stage_in(ctx, c, 5, 0))) return r;
    int64_t total = 0;
    r = hlsmap_view_format_dev(ctx, n, (int *)c[0].d, (int *)c[1].d, (int *)c[2].d, with_sn,
                               (int *)c[3].d, (char *)c[4].d, capacity, ctx->stream, &total);
    if (total_out) *total_out = total;
    if (r != HLSMAP_OK && r != HLSMAP_ERR_CAPACITY) return r;
    CK(cudaMemcpyAsync(offsets, c[3].d, (size_t)(n + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (r == HLSMAP_OK && total > 0)
        CK(cudaMemcpyAsync(chars, c[4].d, (size_t)total, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return r;
}


// ---------------------------------------------------------------------------------------
//  getSegmentList without any allocation
// ---------------------------------------------------------------------------------------
// Who this is for: code that must not touch the CUDA allocator while it runs -- a media
// server's real-time thread with a no-malloc rule, a plugin running inside a host that owns
// all device memory and forbids cudaMalloc from callbacks, or a caller whose own pool already
// holds every byte of HBM.  hlsmap_get_segment_list_dev() grows the context's tile-status
// scratch with cudaMalloc the first time a batch is larger than any before it; this variant
// never does.  Its scratch is the fixed array below, which lives in every GPU context that
// loads the library (about 0.5 MiB per GPU, until the process exits).
//
// Layout, in 8-byte words: [0,5) ScanCounters, [8,12) DeviceResult, [32, 32+kNoAllocMaxTiles)
// one look-back status word per tile.
namespace {
constexpr long long kNoAllocMaxTiles = 65536;
constexpr int kNoAllocHeaderWords = 32;
constexpr int kNoAllocMaxDevices = 64;
static_assert(sizeof(ScanCounters) <= 8 * 8 && sizeof(DeviceResult) <= (kNoAllocHeaderWords - 8) * 8,
              "no-alloc scratch header layout");
}  // namespace
static __device__ __align__(256) unsigned long long g_noalloc_scratch[kNoAllocHeaderWords + kNoAllocMaxTiles];

namespace {
// `g_noalloc_scratch` is only a stand-in name on the host; the device address has to be asked
// for, once per GPU, with that GPU current.  Nothing here allocates.
void *g_noalloc_ptr[kNoAllocMaxDevices] = {};

int noalloc_scratch_for(hlsmap_ctx *ctx, unsigned long long **out)
{
    if (ctx->device < 0 || ctx->device >= kNoAllocMaxDevices)
        return fail(HLSMAP_ERR_INVALID, "no-alloc entry point: device ordinal beyond its table");
    if (!g_noalloc_ptr[ctx->device]) {
        void *ptr = nullptr;
        CK(cudaSetDevice(ctx->device));
        cudaError_t e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr) {
            cudaGetLastError();
            return fail(HLSMAP_ERR_CUDA, std::string("cudaGetSymbolAddress(no-alloc scratch): ") +
                                             cudaGetErrorString(e));
        }
        g_noalloc_ptr[ctx->device] = ptr;
    }
    *out = static_cast<unsigned long long *>(g_noalloc_ptr[ctx->device]);
    return HLSMAP_OK;
}
}  // namespace

int hlsmap_get_segment_list_noalloc_dev(hlsmap_map *map, int64_t nq, const int32_t *d_playlist,
                                        const int32_t *d_level, const double *d_begin,
                                        const double *d_duration, int32_t *d_offsets,
                                        int32_t *d_sn_out, double *d_time_out,
                                        int64_t hit_capacity, uint8_t *d_status, void *stream,
                                        hlsmap_result *result)
{
    if (!map) return fail(HLSMAP_ERR_INVALID, "null map");
    hlsmap_ctx *ctx = map->ctx;
    int r = use_ctx(ctx);
    if (r) return r;
    if (nq < 0 || !d_offsets || (nq > 0 && (!d_level || !d_begin || !d_duration)))
        return fail(HLSMAP_ERR_INVALID, "get_segment_list_noalloc: null query/offset arrays");
    if ((d_sn_out == nullptr) != (d_time_out == nullptr))
        return fail(HLSMAP_ERR_INVALID, "get_segment_list_noalloc: sn_out and time_out go together");
    if (hit_capacity < 0) return fail(HLSMAP_ERR_INVALID, "negative hit capacity");
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;

    const int path = pick_path(map);
    const TilePlan plan = plan_tiles(map, path, nq);
    const int tile_q = plan.tile_q;
    const int64_t n_tiles = plan.n_tiles;
    if (n_tiles > kNoAllocMaxTiles)
        return fail(HLSMAP_ERR_INVALID,
                    "get_segment_list_noalloc: batch of " + std::to_string(nq) +
                        " queries needs " + std::to_string(n_tiles) + " tiles, the fixed scratch holds " +
                        std::to_string(kNoAllocMaxTiles) + " (at most " +
                        std::to_string(kNoAllocMaxTiles * tile_q) +
                        " queries per call on this map): split the batch");

    unsigned long long *scratch = nullptr;
    r = noalloc_scratch_for(ctx, &scratch);   // first call on a GPU: looks the array up
    if (r) return r;
    ScanCounters *ctr = reinterpret_cast<ScanCounters *>(scratch);
    DeviceResult *dres = reinterpret_cast<DeviceResult *>(scratch + 8);
    unsigned long long *tile_status = scratch + kNoAllocHeaderWords;

    // The array keeps whatever the previous call left: the header counters and the status
    // words this batch will read must start at zero (stream-ordered, no allocation).
    CK(cudaMemsetAsync(scratch, 0, (size_t)(kNoAllocHeaderWords + (n_tiles > 0 ? n_tiles : 0)) * 8, s));
    if (nq == 0) {
        CK(cudaMemsetAsync(d_offsets, 0, sizeof(int32_t), s));
    } else {
        ResolveParams p{};
        p.desc = map->d_desc;
        p.playlist_base = map->d_playlist_base;
        p.n_playlists = map->n_playlists;
        p.start = map->d_start;
        p.sn = map->d_sn;
        p.n_frag_padded = map->n_frag_padded;
        p.nq = nq;
        p.q_playlist = d_playlist;
        p.q_level = d_level;
        p.q_begin = d_begin;
        p.q_duration = d_duration;
        p.offsets = d_offsets;
        p.sn_out = d_sn_out;
        p.time_out = d_time_out;
        p.capacity = d_sn_out ? hit_capacity : 0;
        p.status = d_status;
        p.tile_status = tile_status;
        p.ctr = ctr;
        p.result = dres;
        p.epoch = 1;   // words were just zeroed, so any non-zero epoch marks "written by this call"
        p.tslot = 0;
        p.n_tiles = (int)n_tiles;
        p.tile_q = tile_q;
        p.stage_elems = map->stage_elems;
        p.stage_all = map->info.stage_all;
        p.small = map->info.stage_all;
        p.sn_contig = (map->sn_contig && !map->info.stage_all) ? 1 : 0;
        p.uniform_levels = map->uniform_levels;
        p.ipt = map->ipt;
        p.hint = map->d_hint;
        p.grid = map->d_grid;
        p.grid_padded = map->grid_padded;
        p.n_tracks_padded = map->n_tracks_padded;
        p.n_base_padded = map->n_base_padded;
        p.vec_ok = al16(d_level) && al16(d_begin) && al16(d_duration) && al16(d_offsets) &&
                   (!d_playlist || al16(d_playlist)) &&
                   (!d_status || (reinterpret_cast<uintptr_t>(d_status) & 3u) == 0);
        if (path == HLSMAP_PATH_SEARCH) {
            const int grid = (int)(n_tiles < map->grid_search ? n_tiles : map->grid_search);
            CK(launch_resolve_search(p, grid, map->stage_smem, s));
        } else {
            const int grid = (int)(n_tiles < map->grid_filter ? n_tiles : map->grid_filter);
            CK(launch_resolve_filter(p, grid, s));
        }
    }
    if (!result) return HLSMAP_OK;

    // h_result is the context's pinned 32-byte mailbox, allocated when the context was made
    CK(cudaMemcpyAsync(ctx->h_result, dres, sizeof(DeviceResult), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    result->total_hits = ctx->h_result->total_hits;
    result->n_unparsed = ctx->h_result->n_unparsed;
    result->n_missing = ctx->h_result->n_missing;
    result->kernel_launches = nq > 0 ? 1 : 0;
    result->path = path;
    if (ctx->h_result->overflow)
        return fail(HLSMAP_ERR_OVERFLOW, "batch produced more than INT32_MAX hits; split it");
    if (d_sn_out && result->total_hits > hit_capacity)
        return fail(HLSMAP_ERR_CAPACITY, "hit buffers too small: need " +
                                             std::to_string(result->total_hits) + " entries");
    return HLSMAP_OK;
}

}  // extern "C"

/* ---- JSON form of views (json.cu) ---------------------------------------- */
int hlsmap_view_to_json_dev(hlsmap_ctx *ctx, int64_t n, const int32_t *d_level, const int32_t *d_url_id,
                            const int32_t *d_sn, const double *d_time, int32_t *d_offsets, char *d_chars,
                            int64_t capacity, void *stream, int64_t *total_out)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || !d_offsets || capacity < 0 ||
        (n > 0 && (!d_level || !d_url_id || !d_sn || (capacity > 0 && !d_chars))))
        return fail(HLSMAP_ERR_INVALID, "view_to_json: null arrays");
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    ctx->have_last = true;
    ctx->last_path = 0;
    ctx->last_launches = 0;
    if (n == 0) {
        CK(cudaMemsetAsync(d_offsets, 0, sizeof(int32_t), s));
        if (total_out) *total_out = 0;
        return HLSMAP_OK;
    }
    const int64_t n_tiles = (n + kJsonTileViews - 1) / kJsonTileViews;
    if (n_tiles > 0x7fffffff) return fail(HLSMAP_ERR_INVALID, "batch too large: split it");
    void *st_d = nullptr, *st_m = nullptr, *st_b = nullptr;      // grow-only context scratch
    if ((r = ensure_slot(ctx, 8, n * 8, &st_d, s)) || (r = ensure_slot(ctx, 9, n * 4, &st_m, s)) ||
        (r = ensure_slot(ctx, 10, (n_tiles + 1) * 16, &st_b, s)))
        return r;
    JsonWriteParams p{};
    p.n = n;
    p.level = d_level;
    p.url_id = d_url_id;
    p.sn = d_sn;
    p.time = d_time;
    p.offsets = d_offsets;
    p.chars = d_chars;
    p.capacity = capacity;
    p.stage_digits = (unsigned long long *)st_d;
    p.stage_meta = (unsigned *)st_m;
    p.block_sum = (unsigned long long *)st_b;
    p.block_base = (unsigned long long *)st_b + n_tiles + 1;
    p.result = result_of(ctx);
    p.n_tiles = (int)n_tiles;
    const int maxg = view_to_json_max_grid();
    CK(launch_view_to_json(p, (int)(n_tiles < maxg ? n_tiles : maxg), s));
    ctx->last_launches = 3;
    if (total_out) {
        hlsmap_result res{};
        r = fetch_result(ctx, s, &res);
        if (r) return r;
        *total_out = res.total_hits;
        if (res.total_hits > capacity)
            return fail(HLSMAP_ERR_CAPACITY, "character buffer too small: need " +
                                                 std::to_string(res.total_hits) + " bytes");
    }
    return HLSMAP_OK;
}

int hlsmap_view_to_json(hlsmap_ctx *ctx, int64_t n, const int32_t *level, const int32_t *url_id,
                        const int32_t *sn, const double *time, int32_t *offsets, char *chars,
                        int64_t capacity, int64_t *total_out)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || !offsets || capacity < 0 || (n > 0 && (!level || !url_id || !sn || (capacity > 0 && !chars))))
        return fail(HLSMAP_ERR_INVALID, "view_to_json: null arrays");
    const int64_t n1 = n > 0 ? n : 1;
    Cols c[6] = {{level, n1 * 4, nullptr},  {url_id, n1 * 4, nullptr},       {sn, n1 * 4, nullptr},
                 {time, n1 * 8, nullptr},   {nullptr, (n1 + 1) * 4, nullptr}, {nullptr, capacity > 0 ? capacity : 16, nullptr}};
    if (n == 0) c[0].h = c[1].h = c[2].h = c[3].h = nullptr;
    if ((r = stage_in(ctx, c, 6, 0))) return r;
    int64_t total = 0;
    r = hlsmap_view_to_json_dev(ctx, n, (int *)c[0].d, (int *)c[1].d, (int *)c[2].d, time ? (double *)c[3].d : nullptr,
                                (int *)c[4].d, (char *)c[5].d, capacity, ctx->stream, &total);
    if (total_out) *total_out = total;
    if (r != HLSMAP_OK && r != HLSMAP_ERR_CAPACITY) return r;
    const std::string keep = g_err;
    CK(cudaMemcpyAsync(offsets, c[4].d, (size_t)(n + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (r == HLSMAP_OK && total > 0)
        CK(cudaMemcpyAsync(chars, c[5].d, (size_t)total, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (r) g_err = keep;
    return r;
}

int hlsmap_view_from_json_dev(hlsmap_ctx *ctx, int64_t n, const char *d_chars, const int32_t *d_offsets,
                              int32_t *d_level, int32_t *d_url_id, int32_t *d_sn, double *d_time,
                              uint8_t *d_flags, uint8_t *d_status, void *stream)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && (!d_chars || !d_offsets || !d_level || !d_url_id || !d_sn || !d_time || !d_flags ||
                            !d_status)))
        return fail(HLSMAP_ERR_INVALID, "view_from_json: null arrays");
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    JsonReadParams p{};
    p.n = n;
    p.chars = d_chars;
    p.vec_ok = ((uintptr_t)d_chars & 15) == 0;
    p.offsets = d_offsets;
    p.level = d_level;
    p.url_id = d_url_id;
    p.sn = d_sn;
    p.time = d_time;
    p.flags = d_flags;
    p.status = d_status;
    CK(launch_view_from_json(p, s));
    ctx->have_last = true;
    ctx->last_path = 0;
    ctx->last_launches = n > 0 ? 1 : 0;
    return HLSMAP_OK;
}

int hlsmap_view_from_json(hlsmap_ctx *ctx, int64_t n, const char *chars, const int32_t *offsets,
                          int32_t *level, int32_t *url_id, int32_t *sn, double *time, uint8_t *flags,
                          uint8_t *status)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n < 0 || (n > 0 && (!chars || !offsets || !level || !url_id || !sn || !time || !flags || !status)))
        return fail(HLSMAP_ERR_INVALID, "view_from_json: null arrays");
    if (n == 0) return HLSMAP_OK;
    if (offsets[0] != 0) return fail(HLSMAP_ERR_INVALID, "view_from_json: offsets[0] != 0");
    for (int64_t i = 0; i < n; i++)
        if (offsets[i + 1] < offsets[i]) return fail(HLSMAP_ERR_INVALID, "view_from_json: offsets not monotone");
    const int64_t nc = offsets[n];
    Cols c[8] = {{chars, nc > 0 ? nc : 16, nullptr}, {offsets, (n + 1) * 4, nullptr}, {nullptr, n * 4, nullptr},
                 {nullptr, n * 4, nullptr},          {nullptr, n * 4, nullptr},       {nullptr, n * 8, nullptr},
                 {nullptr, n, nullptr},              {nullptr, n, nullptr}};
    if (nc == 0) c[0].h = nullptr;
    if ((r = stage_in(ctx, c, 8, 0))) return r;
    r = hlsmap_view_from_json_dev(ctx, n, (char *)c[0].d, (int *)c[1].d, (int *)c[2].d, (int *)c[3].d, (int *)c[4].d,
                                  (double *)c[5].d, (uint8_t *)c[6].d, (uint8_t *)c[7].d, ctx->stream);
    if (r) return r;
    cudaStream_t s = ctx->stream;
    CK(cudaMemcpyAsync(level, c[2].d, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(url_id, c[3].d, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(sn, c[4].d, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(time, c[5].d, (size_t)n * 8, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(flags, c[6].d, (size_t)n, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(status, c[7].d, (size_t)n, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return HLSMAP_OK;
}

/* ---- media playlists (M3U8 text) -> fragment tables (ingest.cu) ------------- */
namespace {

struct DevBuf {   // frees on scope exit; the ingest entry points are set-up time, not hot
    void *p = nullptr;
    ~DevBuf()
    {
        if (p) cudaFree(p);
    }
    int alloc(size_t bytes)
    {
        if (p) {
            cudaFree(p);
            p = nullptr;
        }
        cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
        if (e != cudaSuccess) {
            cudaGetLastError();
            p = nullptr;
            return fail(HLSMAP_ERR_NOMEM, "cudaMalloc of " + std::to_string(bytes) + " bytes failed");
        }
        return HLSMAP_OK;
    }
};

size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// Everything on the device.  d_start / d_sn / d_dur may be NULL with capacity 0 (count only:
// track_off, live and the totals are still produced).  d_dur0 may be NULL.
int ingest_core(hlsmap_ctx *ctx, int32_t n_tracks, const uint8_t *d_state, const unsigned char *d_text,
                const int64_t *d_text_off, int64_t n_bytes, int64_t *d_track_off, uint8_t *d_live,
                double *d_start, int32_t *d_sn, double *d_dur, double *d_dur0, int64_t capacity,
                cudaStream_t s, hlsmap_ingest_result *res)
{
    if (n_bytes > (int64_t)0x7fffffff * (kIngestSpanBytes / 2))
        return fail(HLSMAP_ERR_INVALID, "parse_m3u8: text too large for one call");
    const int n_tiles = (int)((n_bytes + kIngestTileBytes - 1) / kIngestTileBytes);
    int64_t tag_cap = (int64_t)n_tracks * 2 + 1024;
    IngestCounters h{};
    int launches = 0;
    const bool fill = d_start && d_sn && d_dur;
    if (!fill) capacity = 0;
    for (int attempt = 0; attempt < 2; attempt++) {
        // zeroed head (block sums, counters), then everything the kernels fully overwrite
        const int64_t n_spans = (int64_t)n_tiles * (kIngestTileBytes / kIngestSpanBytes);
        const int64_t n_blocks = (n_spans + kIngestScanBlock - 1) / kIngestScanBlock;
        size_t o_bs = 0, o_out = align256(o_bs + 16 * (size_t)(n_blocks + 1));
        size_t o_zero_end = align256(o_out + sizeof(IngestCounters));
        size_t o_tlo = o_zero_end, o_bf = align256(o_tlo + sizeof(int) * (size_t)(n_tiles + 1));
        size_t o_sc = align256(o_bf + sizeof(int) * (size_t)(n_tiles + 1));
        size_t o_bb = align256(o_sc + 4 * (size_t)(n_spans + 1));
        size_t o_sb = align256(o_bb + 16 * (size_t)(n_blocks + 1));
        size_t o_tp = align256(o_sb + 16 * (size_t)(n_spans + 1));
        size_t o_tf = align256(o_tp + 8 * (size_t)tag_cap);
        size_t o_tv = align256(o_tf + 8 * (size_t)tag_cap);
        size_t o_d0 = align256(o_tv + 4 * (size_t)tag_cap);
        size_t o_sd = align256(o_d0 + 8 * (size_t)(n_tracks + 1));
        size_t o_sg = align256(o_sd + (fill ? 8 * (size_t)kIngestSpanFrags * (size_t)n_spans : 0));
        size_t total = align256(o_sg + (fill ? 8 * (size_t)kIngestSpanTags * (size_t)n_spans : 0));
        void *sp = nullptr;               // grow-only context scratch: no allocation after the first call
        int r = ensure_slot(ctx, kSlots - 1, (int64_t)total, &sp, s);
        if (r) return r;
        char *b = (char *)sp;
        CK(cudaMemsetAsync(b, 0, o_zero_end, s));
        if (n_tracks) CK(cudaMemsetAsync(d_live, 1, (size_t)n_tracks, s));
        CK(cudaMemsetAsync(d_track_off, 0, 8 * (size_t)(n_tracks + 1), s));
        IngestParams p{};
        p.text = d_text;
        p.n_bytes = n_bytes;
        p.text_off = (const long long *)d_text_off;
        p.n_tracks = n_tracks;
        p.state = d_state;
        p.tile_tlo = (int *)(b + o_tlo);
        p.tile_bfirst = (int *)(b + o_bf);
        p.dur = d_dur;
        p.capacity = capacity;
        p.track_off = (long long *)d_track_off;
        p.live = d_live;
        p.tag_pos = (long long *)(b + o_tp);
        p.tag_frag = (long long *)(b + o_tf);
        p.tag_val = (int *)(b + o_tv);
        p.tag_capacity = tag_cap;
        p.stage_dur = fill ? (double *)(b + o_sd) : nullptr;
        p.stage_tag = fill ? (unsigned long long *)(b + o_sg) : nullptr;
        p.span_cnt = (unsigned *)(b + o_sc);
        p.block_sum = (unsigned long long *)(b + o_bs);
        p.block_base = (unsigned long long *)(b + o_bb);
        p.span_base = (unsigned long long *)(b + o_sb);
        p.n_spans = (int)n_spans;
        p.n_blocks = (int)n_blocks;
        p.n_tiles = n_tiles;
        p.vec_ok = ((uintptr_t)d_text & 15) == 0;
        p.out = (IngestCounters *)(b + o_out);
        static const bool timing = getenv("HLSMAP_INGEST_TIMING") != nullptr;     // diagnostics
        cudaEvent_t ev[4] = {};
        if (timing)
            for (auto &e : ev) cudaEventCreate(&e);
        if (timing) cudaEventRecord(ev[0], s);
        if (n_tiles > 0 && n_tracks > 0) {
            CK(launch_ingest_tile_index(p.text_off, n_tracks, n_tiles, (int *)(b + o_tlo), (int *)(b + o_bf), s));
            if (timing) cudaEventRecord(ev[1], s);
            CK(launch_ingest_scan(p, s));
            CK(launch_ingest_compact(p, s));
            launches += fill ? 6 : 5;
        }
        if (timing) cudaEventRecord(ev[2], s);
        // the per-track assembly follows without a host round trip: it reads the counters itself
        // and steps aside when the columns or the tag list are too small
        if (fill && n_tracks > 0) {
            AssembleParams a{};
            a.n_tracks = n_tracks;
            a.state = d_state;
            a.text_off = p.text_off;
            a.track_off = p.track_off;
            a.dur = d_dur;
            a.tag_pos = p.tag_pos;
            a.tag_frag = p.tag_frag;
            a.tag_val = p.tag_val;
            a.capacity = capacity;
            a.tag_capacity = tag_cap;
            a.start = d_start;
            a.sn = d_sn;
            a.dur0 = d_dur0 ? d_dur0 : (double *)(b + o_d0);
            a.out = p.out;
            CK(launch_ingest_assemble(a, n_bytes >= (int64_t)n_tracks * 768, s));   // long tracks: a warp each
            launches++;
        }
        if (timing) cudaEventRecord(ev[3], s);
        CK(cudaMemcpyAsync(&h, p.out, sizeof h, cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        if (timing && n_tiles > 0) {
            float t_idx = 0, t_scan = 0, t_asm = 0;
            cudaEventElapsedTime(&t_idx, ev[0], ev[1]);
            cudaEventElapsedTime(&t_scan, ev[1], ev[2]);
            cudaEventElapsedTime(&t_asm, ev[2], ev[3]);
            fprintf(stderr, "[hlsmap ingest] tile index %.1f us, scan+compact %.1f us, assemble %.1f us\n", 1e3 * t_idx,
                    1e3 * t_scan, 1e3 * t_asm);
        }
        if ((int64_t)h.n_tags > tag_cap) {     // more sequence tags than tracks*2: size exactly, redo
            tag_cap = (int64_t)h.n_tags;
            if (timing)
                for (auto &e : ev) cudaEventDestroy(e);
            continue;
        }
        if (timing)
            for (auto &e : ev) cudaEventDestroy(e);
        break;
    }
    if (res) {
        res->n_fragments = (int64_t)h.n_fragments;
        res->n_sequence_tags = (int64_t)h.n_tags;
        res->n_unsupported = (int64_t)h.n_unsupported;
        res->text_bytes = n_bytes;
        res->kernel_launches = launches;
        res->reserved = 0;
    }
    if (h.n_unsupported)
        return fail(HLSMAP_ERR_UNSUPPORTED,
                    "parse_m3u8: " + std::to_string(h.n_unsupported) +
                        " item(s) the parser refuses rather than guesses: a number it cannot convert exactly "
                        "(more than 19 significant digits with an undecidable tail, a sequence number beyond "
                        "int32), or a tag more than 16384 bytes into its line");
    if (fill && (int64_t)h.n_fragments > capacity)
        return fail(HLSMAP_ERR_CAPACITY, "fragment buffers too small: need " + std::to_string(h.n_fragments) + " entries");
    return HLSMAP_OK;
}

int check_text_offsets(int32_t n_tracks, const int64_t *text_offset)
{
    if (n_tracks < 0 || !text_offset) return fail(HLSMAP_ERR_INVALID, "parse_m3u8: missing text_offset");
    if (text_offset[0] != 0) return fail(HLSMAP_ERR_INVALID, "parse_m3u8: text_offset[0] != 0");
    for (int32_t t = 0; t < n_tracks; t++)
        if (text_offset[t + 1] < text_offset[t]) return fail(HLSMAP_ERR_INVALID, "parse_m3u8: text_offset not monotone");
    return HLSMAP_OK;
}

}  // namespace

int hlsmap_parse_m3u8_dev(hlsmap_ctx *ctx, int32_t n_tracks, const uint8_t *track_state, const char *text,
                          const int64_t *text_offset, int64_t text_bytes, int64_t *track_offset,
                          uint8_t *track_live, double *start, int32_t *sn, double *duration,
                          int64_t capacity, hlsmap_ingest_result *res, void *stream)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (n_tracks < 0 || text_bytes < 0 || !text_offset || !track_offset || (n_tracks > 0 && !track_live) ||
        (text_bytes > 0 && !text))
        return fail(HLSMAP_ERR_INVALID, "parse_m3u8_dev: missing arrays");
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    return ingest_core(ctx, n_tracks, track_state, (const unsigned char *)text, text_offset, text_bytes,
                       track_offset, track_live, start, sn, duration, nullptr, capacity, s, res);
}

int hlsmap_parse_m3u8(hlsmap_ctx *ctx, int32_t n_tracks, const uint8_t *track_state, const char *text,
                      const int64_t *text_offset, int64_t *track_offset, uint8_t *track_live, double *start,
                      int32_t *sn, double *duration, int64_t capacity, hlsmap_ingest_result *res)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if ((r = check_text_offsets(n_tracks, text_offset))) return r;
    if (!track_offset || (n_tracks > 0 && !track_live))
        return fail(HLSMAP_ERR_INVALID, "parse_m3u8: missing output arrays");
    const int64_t nb = text_offset[n_tracks];
    if (nb > 0 && !text) return fail(HLSMAP_ERR_INVALID, "parse_m3u8: null text");
    const bool fill = start && sn && duration && capacity > 0;
    if (!fill) capacity = 0;
    cudaStream_t s = ctx->stream;
    DevBuf d_text, d_off, d_state, d_toff, d_live, d_start, d_sn, d_dur;
    if ((r = d_text.alloc((size_t)nb + 16)) || (r = d_off.alloc(8 * (size_t)(n_tracks + 1))) ||
        (r = d_state.alloc((size_t)n_tracks + 1)) || (r = d_toff.alloc(8 * (size_t)(n_tracks + 1))) ||
        (r = d_live.alloc((size_t)n_tracks + 1)))
        return r;
    if (fill && ((r = d_start.alloc(8 * (size_t)capacity)) || (r = d_sn.alloc(4 * (size_t)capacity)) ||
                 (r = d_dur.alloc(8 * (size_t)capacity))))
        return r;
    if (nb) CK(cudaMemcpyAsync(d_text.p, text, (size_t)nb, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_off.p, text_offset, 8 * (size_t)(n_tracks + 1), cudaMemcpyHostToDevice, s));
    if (track_state && n_tracks)
        CK(cudaMemcpyAsync(d_state.p, track_state, (size_t)n_tracks, cudaMemcpyHostToDevice, s));
    hlsmap_ingest_result local{};
    r = ingest_core(ctx, n_tracks, track_state ? (const uint8_t *)d_state.p : nullptr,
                    (const unsigned char *)d_text.p, (const int64_t *)d_off.p, nb, (int64_t *)d_toff.p,
                    (uint8_t *)d_live.p, (double *)d_start.p, (int32_t *)d_sn.p, (double *)d_dur.p, nullptr,
                    capacity, s, &local);
    if (res) *res = local;
    if (r != HLSMAP_OK && r != HLSMAP_ERR_CAPACITY) return r;
    const std::string keep = g_err;
    CK(cudaMemcpyAsync(track_offset, d_toff.p, 8 * (size_t)(n_tracks + 1), cudaMemcpyDeviceToHost, s));
    if (n_tracks) CK(cudaMemcpyAsync(track_live, d_live.p, (size_t)n_tracks, cudaMemcpyDeviceToHost, s));
    if (r == HLSMAP_OK && fill && local.n_fragments) {
        const size_t nf = (size_t)local.n_fragments;
        CK(cudaMemcpyAsync(start, d_start.p, 8 * nf, cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(sn, d_sn.p, 4 * nf, cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(duration, d_dur.p, 8 * nf, cudaMemcpyDeviceToHost, s));
    }
    CK(cudaStreamSynchronize(s));
    if (r) g_err = keep;
    return r;
}

// Parse text that is already on the device (d_text, d_off) and build the map from the columns.
static int map_from_device_text(hlsmap_ctx *ctx, int32_t n_playlists, const int32_t *playlist_track_base,
                                int32_t n_tracks, const uint8_t *track_state, const int32_t *url_count,
                                const unsigned char *d_text, const int64_t *d_off, int64_t nb,
                                hlsmap_ingest_result *res, hlsmap_map **out)
{
    int r;
    cudaStream_t s = ctx->stream;
    DevBuf d_state, d_toff, d_live, d_start, d_sn, d_dur, d_dur0;
    if ((r = d_state.alloc((size_t)n_tracks + 1)) || (r = d_toff.alloc(8 * (size_t)(n_tracks + 1))) ||
        (r = d_live.alloc((size_t)n_tracks + 1)) || (r = d_dur0.alloc(8 * (size_t)(n_tracks + 1))))
        return r;
    if (n_tracks) CK(cudaMemcpyAsync(d_state.p, track_state, (size_t)n_tracks, cudaMemcpyHostToDevice, s));
    // pass 1 counts (the text is read once more, from L2 when it fits), so the columns are
    // allocated at their exact size instead of the worst case of one fragment per 9 bytes
    hlsmap_ingest_result local{};
    r = ingest_core(ctx, n_tracks, (const uint8_t *)d_state.p, d_text, d_off, nb, (int64_t *)d_toff.p,
                    (uint8_t *)d_live.p, nullptr, nullptr, nullptr, nullptr, 0, s, &local);
    if (res) *res = local;
    if (r) return r;
    const int64_t nf = local.n_fragments;
    if ((r = d_start.alloc(8 * (size_t)nf + 16)) || (r = d_sn.alloc(4 * (size_t)nf + 16)) ||
        (r = d_dur.alloc(8 * (size_t)nf + 16)))
        return r;
    const int first_launches = local.kernel_launches;
    r = ingest_core(ctx, n_tracks, (const uint8_t *)d_state.p, d_text, d_off, nb, (int64_t *)d_toff.p,
                    (uint8_t *)d_live.p, (double *)d_start.p, (int32_t *)d_sn.p, (double *)d_dur.p,
                    (double *)d_dur0.p, nf > 0 ? nf : 1, s, &local);
    local.kernel_launches += first_launches;
    if (res) *res = local;
    if (r) return r;
    std::vector<int64_t> h_off((size_t)n_tracks + 1, 0);
    CK(cudaMemcpyAsync(h_off.data(), d_toff.p, 8 * (size_t)(n_tracks + 1), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    // tracks that are not parsed hold no fragments whatever their text said
    return map_create_impl(ctx, n_playlists, playlist_track_base, n_tracks, h_off.data(), track_state, url_count,
                           (const double *)d_start.p, (const int32_t *)d_sn.p, nullptr, (const double *)d_dur0.p,
                           out);
}

int hlsmap_map_create_from_m3u8(hlsmap_ctx *ctx, int32_t n_playlists, const int32_t *playlist_track_base,
                                int32_t n_tracks, const uint8_t *track_state, const int32_t *url_count,
                                const char *text, const int64_t *text_offset, hlsmap_ingest_result *res,
                                hlsmap_map **out)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (!out) return fail(HLSMAP_ERR_INVALID, "null out pointer");
    *out = nullptr;
    if ((r = check_text_offsets(n_tracks, text_offset))) return r;
    if (n_tracks > 0 && !track_state) return fail(HLSMAP_ERR_INVALID, "map_create_from_m3u8: null track_state");
    const int64_t nb = text_offset[n_tracks];
    if (nb > 0 && !text) return fail(HLSMAP_ERR_INVALID, "map_create_from_m3u8: null text");
    cudaStream_t s = ctx->stream;
    DevBuf d_text, d_off;
    if ((r = d_text.alloc((size_t)nb + 16)) || (r = d_off.alloc(8 * (size_t)(n_tracks + 1)))) return r;
    if (nb) CK(cudaMemcpyAsync(d_text.p, text, (size_t)nb, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(d_off.p, text_offset, 8 * (size_t)(n_tracks + 1), cudaMemcpyHostToDevice, s));
    return map_from_device_text(ctx, n_playlists, playlist_track_base, n_tracks, track_state, url_count,
                                (const unsigned char *)d_text.p, (const int64_t *)d_off.p, nb, res, out);
}

int hlsmap_map_create_from_m3u8_dev(hlsmap_ctx *ctx, int32_t n_playlists, const int32_t *playlist_track_base,
                                    int32_t n_tracks, const uint8_t *track_state, const int32_t *url_count,
                                    const char *d_text, const int64_t *d_text_offset, int64_t text_bytes,
                                    hlsmap_ingest_result *res, hlsmap_map **out)
{
    int r = use_ctx(ctx);
    if (r) return r;
    if (!out) return fail(HLSMAP_ERR_INVALID, "null out pointer");
    *out = nullptr;
    if (n_tracks < 0 || text_bytes < 0 || !d_text_offset || (text_bytes > 0 && !d_text) || (n_tracks > 0 && !track_state))
        return fail(HLSMAP_ERR_INVALID, "map_create_from_m3u8_dev: missing arrays");
    CK(cudaStreamSynchronize(ctx->stream));
    return map_from_device_text(ctx, n_playlists, playlist_track_base, n_tracks, track_state, url_count,
                                (const unsigned char *)d_text, d_text_offset, text_bytes, res, out);
}

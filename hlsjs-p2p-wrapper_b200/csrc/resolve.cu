// resolve.cu -- batched MediaMap.getSegmentList on sm_100a.
//
// Reference semantics (lib/integration/mapping/media-map.js:27-54): for one query keep,
// in index order, every fragment with
//        beginTime <= start && start <= fl(beginTime + duration)          (:44)
// and emit {sn: fragment.sn, time: fragment.start} per kept fragment (:45-49).
//
// Two kernels produce bit-identical output:
//   resolve_search_kernel  -- when every start column is monotone and NaN-free the kept
//       fragments are the contiguous run [lower_bound(begin), upper_bound(ub)); both
//       bounds are found with the SAME two fp64 comparisons the filter uses, so NaN,
//       +-Inf, -0 and duplicates behave identically.
//   resolve_filter_kernel  -- the reference's linear filter itself (warp per query,
//       ballot/popc ordered compaction) for tables that are not provably sorted.
// Both are single-pass: count -> CTA scan -> decoupled look-back (scan.cuh) -> fill, so
// queries are read once and hits written once.
#include <cfloat>
#include <type_traits>

#include "scan.cuh"

namespace hlsmap {

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kCntBits = 44;                    // fused scan: [nz:20][count:44]
constexpr unsigned long long kCntMask = (1ull << kCntBits) - 1ull;

using Scan = TileScan<kThreads>;

// Output slots one warp writes per round (4 per lane, 32 apart): chunks start at multiples of kChunk
// slots of the OUTPUT arrays, so every store instruction of a warp covers whole 128-byte lines.
constexpr int kChunk = 128;

// Compile-time shape of one search-kernel variant.
//   IPT     queries per thread (tile = 256 * IPT queries)
//   WINDOW  output slots expanded per pass over the tile's output run
template <int IPT>
struct Shape {
    static constexpr int kIpt = IPT;
    static constexpr int kTile = kThreads * IPT;
    // the bitmap reaches one chunk past the window: a chunk that starts inside a window ends inside its bitmap
    static constexpr int kBits = 8192;
    static constexpr int kWindow = kBits - kChunk;
    static constexpr int kWords = kBits / 32;   // <= kThreads: one bitmap word per thread
    static_assert(kWords <= kThreads, "one bitmap word per thread at most");
};

// Expansion records: one per query that kept >= 1 fragment, in query order.  With rel = the
// query's first output slot relative to the tile's output run and src = its first kept fragment,
// output slot q of that query comes from fragment src + (q - rel): the record keeps src - rel.
// SMALL maps keep it as a shared-memory BYTE address: &s_sn[src - rel]; slot q's sn then sits at
// that address + 4*q and its start at 2 * (that address + 4*q) + (addr(s_start) - 2 * addr(s_sn)):
// one add per slot, one shift-add more for the time.
template <bool SMALL, int TILE>
struct RecTable;
template <int TILE>
struct RecTable<true, TILE> {   // every fragment index fits an int
    unsigned v[TILE];
    static constexpr int kShift = 2;
    __device__ __forceinline__ void set(int i, int rel, long long src, unsigned sn_addr, int = 0)
    {
        v[i] = sn_addr + 4u * (unsigned)((int)src - rel);
    }
};
template <int TILE>
struct RecTable<false, TILE> {
    // v = first kept fragment - first slot; sb = sn of the first kept fragment - first slot, for
    // tracks whose sn column is proven contiguous: slot q's sn is then sb + q
    long long v[TILE];
    int sb[TILE];
    static constexpr int kShift = 3;
    __device__ __forceinline__ void set(int i, int rel, long long src, unsigned, int sn_first = 0)
    {
        v[i] = src - rel;
        sb[i] = sn_first - rel;
    }
};

// Shared state of one CTA.
template <bool SMALL, int IPT>
struct SearchSmem {
    Scan::Smem scan;
    unsigned long long mbar;
    long long red_min[2][kWarps];     // double buffered: staging passes of one tile alternate
    long long red_max[2][kWarps];
    unsigned long long warp_cnt[kWarps];
    int warp_heads[kWarps];
    int tile_slot[2];              // next tile id, double buffered across iterations
    unsigned long long tile_base;  // exclusive prefix of the tile being emitted (from warp 0)
    RecTable<SMALL, Shape<IPT>::kTile> rec;
    // per 32 output slots of the window: .x = bitmap of slots where a record starts,
    // .y = records that start before this word (tile-wide rank base).  Double buffered.
    uint2 hb[2][Shape<IPT>::kWords];
    int carried;                      // records that start before the next window
};

__device__ __forceinline__ unsigned smem_u32(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}

// Loads from a 32-bit shared-memory address plus a compile-time offset (one LDS, no address math).
template <int IMM>
__device__ __forceinline__ unsigned lds_u32(unsigned addr)
{
    unsigned v;
    asm("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(IMM));
    return v;
}
template <int IMM>
__device__ __forceinline__ double lds_f64(unsigned addr)
{
    double v;
    asm("ld.shared.f64 %0, [%1+%2];" : "=d"(v) : "r"(addr), "n"(IMM));
    return v;
}
template <int IMM>
__device__ __forceinline__ long long lds_s64(unsigned addr)
{
    long long v;
    asm("ld.shared.s64 %0, [%1+%2];" : "=l"(v) : "r"(addr), "n"(IMM));
    return v;
}

// ---- TMA (bulk async copy) staging of table columns ------------------------------------
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes,
                                         unsigned long long *bar)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void bulk_g2s_chunked(unsigned char *dst, const void *src,
                                                 unsigned bytes, unsigned long long *bar)
{
    for (unsigned o = 0; o < bytes; o += 32768u)
        bulk_g2s(dst + o, reinterpret_cast<const unsigned char *>(src) + o,
                 min(32768u, bytes - o), bar);
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// ---- level lookup: media-map.js:28-37 --------------------------------------------------
// IN_SMEM: the descriptor arrays were staged in shared memory (plain loads instead of __ldg).
// uniform_levels > 0: every playlist of the map has exactly that many levels (established at upload),
// so its first track is playlist * uniform_levels and the playlist index need not be read: one
// dependent memory round trip less per query.
template <bool IN_SMEM = false>
__device__ __forceinline__ int lookup_track(const TrackDesc *desc, const int *playlist_base,
                                            int n_playlists, int playlist, int level,
                                            long long &off, int &len, unsigned &flags,
                                            int *track = nullptr, int uniform_levels = 0)
{
    off = 0;
    len = 0;
    flags = 0;
    if (track) *track = 0;
    if ((unsigned)playlist >= (unsigned)n_playlists) return HLSMAP_ST_MISSING;
    int b, e;
    if (!IN_SMEM && uniform_levels > 0) {
        b = playlist * uniform_levels;
        e = b + uniform_levels;
    } else {
        b = IN_SMEM ? playlist_base[playlist] : __ldg(playlist_base + playlist);
        e = IN_SMEM ? playlist_base[playlist + 1] : __ldg(playlist_base + playlist + 1);
    }
    if ((unsigned)level >= (unsigned)(e - b)) return HLSMAP_ST_MISSING;  // levels[level] undefined
    const int4 raw = IN_SMEM ? *reinterpret_cast<const int4 *>(desc + b + level)
                             : __ldg(reinterpret_cast<const int4 *>(desc + b + level));
    if (track) *track = b + level;
    flags = (unsigned)raw.w;
    const unsigned st = flags & kTrackStateMask;
    if (st == HLSMAP_TRACK_MISSING) return HLSMAP_ST_MISSING;    // falsy level (:30)
    if (st == HLSMAP_TRACK_UNPARSED) return HLSMAP_ST_UNPARSED;  // no details (:34)
    off = ((long long)(unsigned)raw.x) | ((long long)raw.y << 32);
    len = raw.z;
    return HLSMAP_ST_OK;
}

struct SmemCol {  // start column staged in shared memory; index relative to the span
    const double *s;
    __device__ __forceinline__ double operator()(int i) const { return s[i]; }
};
struct GmemCol {
    const double *g;
    __device__ __forceinline__ double operator()(int i) const { return __ldg(g + i); }
};

// First i in [lo, lo+n) with  begin <= col(i)  (the filter's left test); lo+n if none.
template <class Col>
__device__ __forceinline__ int lower_bound_begin(const Col &col, int lo, int n, double begin)
{
    while (n > 0) {
        const int half = n >> 1;
        const bool keep_right = !(begin <= col(lo + half));
        lo = keep_right ? lo + half + 1 : lo;
        n = keep_right ? n - half - 1 : half;
    }
    return lo;
}

// First i in [lo, lo+n) with  !(col(i) <= ub)  (the filter's right test fails).
template <class Col>
__device__ __forceinline__ int upper_bound_ub(const Col &col, int lo, int n, double ub)
{
    while (n > 0) {
        const int half = n >> 1;
        const bool keep_right = col(lo + half) <= ub;
        lo = keep_right ? lo + half + 1 : lo;
        n = keep_right ? n - half - 1 : half;
    }
    return lo;
}

// Same bound, galloping from lo: a look-ahead window keeps only a handful of fragments.
template <class Col>
__device__ __forceinline__ int upper_bound_from(const Col &col, int lo, int len, double ub)
{
    int a = lo, b = len, step = 1;
    for (;;) {
        const int pr = a + step - 1;
        if (pr >= len) break;
        if (col(pr) <= ub) {
            a = pr + 1;
            step <<= 1;
        } else {
            b = pr;
            break;
        }
    }
    return upper_bound_ub(col, a, b - a, ub);
}

// Both bounds of K queries at once: 2*K independent halving chains, the same number of steps for
// every lane of the warp (the longest track decides), no branches -- so the probes of all chains
// are in flight together.  Query k searches col[k][0, n[k]) (n[k] = 0: not taking part).  On a
// sorted NaN-free column `start <= ub` holds on a prefix, so its end found over the whole track
// equals the one found from the lower bound on.  Every lane of the warp must call this.
template <int K, bool GLOBAL>
__device__ __forceinline__ void bounds_interleaved(const double *const (&col)[K], const int (&n)[K],
                                                   const double (&bg)[K], const double (&ub)[K],
                                                   int (&l)[K], int (&h)[K])
{
    int nl[K], nh[K], mlen = 0;
#pragma unroll
    for (int k = 0; k < K; k++) {
        l[k] = 0;
        h[k] = 0;
        nl[k] = n[k];
        nh[k] = n[k];
        mlen = max(mlen, n[k]);
    }
    mlen = __reduce_max_sync(0xffffffffu, mlen);
    for (int rem = mlen; rem > 0; rem >>= 1) {
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int hl = nl[k] >> 1, hh = nh[k] >> 1;
            // a finished chain keeps probing element 0 (its answer is ignored)
            const double *pl = col[k] + (nl[k] > 0 ? l[k] + hl : 0);
            const double *ph = col[k] + (nh[k] > 0 ? h[k] + hh : 0);
            const double xl = GLOBAL ? __ldg(pl) : *pl;
            const double xh = GLOBAL ? __ldg(ph) : *ph;
            const bool gl = nl[k] > 0 && !(bg[k] <= xl);
            const bool gh = nh[k] > 0 && (xh <= ub[k]);
            l[k] = gl ? l[k] + hl + 1 : l[k];
            nl[k] = gl ? nl[k] - hl - 1 : hl;
            h[k] = gh ? h[k] + hh + 1 : h[k];
            nh[k] = gh ? nh[k] - hh - 1 : hh;
        }
    }
}

__device__ __forceinline__ int warp_excl_scan(int v, int &total)
{
    const int lane = threadIdx.x & 31;
    int x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
    }
    total = __shfl_sync(0xffffffffu, x, 31);
    return x - v;
}

__device__ __forceinline__ unsigned long long warp_excl_scan64(unsigned long long v,
                                                               unsigned long long &total)
{
    const int lane = threadIdx.x & 31;
    unsigned long long x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
    }
    total = __shfl_sync(0xffffffffu, x, 31);
    return x - v;
}

}  // namespace

// Per-tile phase timestamps: diagnostics, compiled in only with -DHLSMAP_TRACE_BUILD=1
// (scripts/trace_run.py builds such a library); the product build carries none of it.
#ifndef HLSMAP_TRACE_BUILD
#define HLSMAP_TRACE_BUILD 0
#endif
#ifndef HLSMAP_WIDE_CTAS
#define HLSMAP_WIDE_CTAS 3        // CTAs per SM the 1024-query variant of the shared-memory kernel is compiled for (4: 64 registers, 36 bytes of spills; 33.5 against 33.1 us per 10^6-query batch, 27.1 against 23.6 on the look-ahead class: 977 tiles over 592 CTAs)
#endif
#ifndef HLSMAP_GENERAL_CTAS
#define HLSMAP_GENERAL_CTAS 3     // CTAs per SM the general kernel is compiled for (4: 64 registers, spills; measured slower)
#endif
__device__ __forceinline__ void trace_mark(const ResolveParams &p, int tile, int slot, int who = 0)
{
#if HLSMAP_TRACE_BUILD
    if (p.trace != nullptr && threadIdx.x == who) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.trace[(long long)tile * 16 + slot] = t;
        if (slot == 0) {
            unsigned smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            p.trace[(long long)tile * 16 + 15] = smid;
        }
    }
#endif
}

// Same, but issued only once `dep` has been produced (e.g. a loaded value).
__device__ __forceinline__ void trace_mark_after(const ResolveParams &p, int tile, int slot, int dep,
                                                 int who = 0)
{
#if HLSMAP_TRACE_BUILD
    if (p.trace != nullptr && threadIdx.x == who) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) : "r"(dep));
        p.trace[(long long)tile * 16 + slot] = t;
    }
#endif
}

// Hint-grid cell of a time value: monotone non-decreasing in x, NaN -> 0, saturating.  The
// SAME function places table entries (upload) and query bounds (lookup), so for a sorted
// NaN-free column  cell(s) < cell(x) => s < x  and  cell(s) > cell(x) => s > x : a bound
// always lies inside [grid[cell], grid[cell+1]] and is finished there with the filter's own
// comparison.  The grid narrows the search; it never decides a comparison.
__device__ __forceinline__ int hint_cell(double x, double t0, double inv, int ncell)
{
    const int c = __double2int_rd(__dmul_rn(__dsub_rn(x, t0), inv));
    return min(max(c, 0), ncell - 1);
}

// Bounds of K queries on a map staged in shared memory (SMALL).  Hint grid: each bound is
// narrowed to one cell before any table probe; the per-query chains are written side by side so
// they overlap (ILP), and the cell is finished with the filter's own comparisons
// (media-map.js:44): [lo, lo + cnt) is exactly what the reference loop keeps.
template <int K>
__device__ __forceinline__ void small_bounds(const double *s_start, const unsigned short *s_grid,
                                             const TrackHint *s_hint, const double (&bg)[K],
                                             const double (&du)[K], const long long (&off)[K],
                                             const int (&len)[K], const int (&st)[K], const int (&trk)[K],
                                             int (&lo)[K], int (&cnt)[K])
{
    int b0[K], bn[K], u0[K], un[K];
    double ub[K];
#pragma unroll
    for (int k = 0; k < K; k++) {
        // one fp64 add, round-to-nearest, never contracted: `beginTime + duration`
        ub[k] = __dadd_rn(bg[k], du[k]);
        const int4 h0 = *reinterpret_cast<const int4 *>(s_hint + trk[k]);
        const int4 h1 = *(reinterpret_cast<const int4 *>(s_hint + trk[k]) + 1);
        const double t0 = __hiloint2double(h0.y, h0.x);
        const double inv = __hiloint2double(h0.w, h0.z);
        const bool ok = st[k] == HLSMAP_ST_OK;
        const int ncell = ok ? h1.y : 1;
        b0[k] = (ok ? h1.x : 0) + hint_cell(bg[k], t0, inv, ncell);   // grid index
        u0[k] = (ok ? h1.x : 0) + hint_cell(ub[k], t0, inv, ncell);
    }
#pragma unroll
    for (int k = 0; k < K; k++) {
        const int gb0 = s_grid[b0[k]], gb1 = s_grid[b0[k] + 1];
        const int gu0 = s_grid[u0[k]], gu1 = s_grid[u0[k] + 1];
        b0[k] = gb0; bn[k] = gb1 - gb0;
        u0[k] = gu0; un[k] = gu1 - gu0;
    }
    // two fixed halving steps settle cells of up to 3 fragments without a loop
#pragma unroll
    for (int it = 0; it < 2; it++) {
#pragma unroll
        for (int k = 0; k < K; k++) {
            const double *col = s_start + (int)off[k];
            const int hb_ = bn[k] >> 1, hu_ = un[k] >> 1;
            const bool gb = bn[k] > 0 && !(bg[k] <= col[b0[k] + hb_]);
            const bool gu = un[k] > 0 && (col[u0[k] + hu_] <= ub[k]);
            b0[k] = gb ? b0[k] + hb_ + 1 : b0[k];
            bn[k] = gb ? bn[k] - hb_ - 1 : hb_;
            u0[k] = gu ? u0[k] + hu_ + 1 : u0[k];
            un[k] = gu ? un[k] - hu_ - 1 : hu_;
        }
    }
#pragma unroll
    for (int k = 0; k < K; k++) {
        const SmemCol col{s_start + (int)off[k]};
        int l = b0[k], h = u0[k];
        if (bn[k] > 0) l = lower_bound_begin(col, l, bn[k], bg[k]);   // crowded cell
        if (un[k] > 0) h = upper_bound_ub(col, h, un[k], ub[k]);
        if (bg[k] != bg[k]) l = len[k];       // NaN: `begin <= start` never holds
        if (st[k] != HLSMAP_ST_OK) { l = 0; h = 0; }
        lo[k] = l;
        cnt[k] = max(h - l, 0);
    }
}

// =========================================================================================
//  search kernel.  SMALL = the whole map (start, sn, hint grid) lives in shared memory.
//
//  Software pipeline, one tile deep: an iteration first SEARCHES tile B (bounds, scan, publish
//  the tile's hit count), then EMITS the tile A it searched one iteration earlier (look-back
//  for A's batch prefix, offsets, hits).  A's predecessors therefore had a whole search phase
//  to publish before A asks for them, so the look-back practically never waits, and B's
//  aggregate is visible to later tiles before A's long emission starts.
// =========================================================================================
// SNC (general kernel only): every parsed track's sn column is proven contiguous (sn[i] = sn[0] + i,
// established on the GPU at upload), so a hit's sn is computed from its query's first hit instead of
// being gathered: one L2 read per query with hits in place of one per hit, and sn[] is not streamed.
template <bool SMALL, int IPT, bool SNC, bool FUSED>
__global__ void __launch_bounds__(kThreads, (SMALL && IPT == 2) ? 4 : (SMALL ? HLSMAP_WIDE_CTAS : (IPT == 2 ? HLSMAP_GENERAL_CTAS : 2)))
resolve_search_kernel(const __grid_constant__ SearchParams<FUSED ? kMaxFused : 1> p)
{
    static_assert(!(SMALL && SNC), "the shared-memory-resident kernel gathers sn from shared memory");
    using S = Shape<IPT>;
    const int nb = FUSED ? p.n_batches : 1;      // batches in this launch
    constexpr int kClaimSlots = 8 * kThreads;   // the next ticket is drawn this many output slots before the tile's last
    constexpr int kIpt = S::kIpt, kWindow = S::kWindow, kWords = S::kWords;
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ SearchSmem<SMALL, IPT> sm;
    double *s_start = reinterpret_cast<double *>(dyn_smem);
    const int *s_sn = reinterpret_cast<const int *>(dyn_smem + p.n_frag_padded * 8);
    const unsigned short *s_grid =
        reinterpret_cast<const unsigned short *>(dyn_smem + p.n_frag_padded * 12);
    // SMALL: track descriptors, hints and the playlist index live in shared memory too (the
    // carve-out leaves almost no L1, so global lookups would each cost an L2 round trip)
    const unsigned char *s_meta = dyn_smem + p.n_frag_padded * 12 + (long long)p.grid_padded * 2;
    const TrackDesc *s_desc = reinterpret_cast<const TrackDesc *>(s_meta);
    const TrackHint *s_hint = reinterpret_cast<const TrackHint *>(s_meta + (long long)p.n_tracks_padded * 16);
    const int *s_base = reinterpret_cast<const int *>(s_meta + (long long)p.n_tracks_padded * 48);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) mbar_init(&sm.mbar, 1);
    if (tid < kWords) {
        sm.hb[0][tid] = make_uint2(0u, 0u);
        sm.hb[1][tid] = make_uint2(0u, 0u);
    }
    unsigned mbar_phase = 0;
    unsigned hbuf = 0;                   // which hb[] buffer the next window uses
    long long span_lo = 0, span_hi = 0;  // staged element range of start[] (general kernel)

    bool stage_pending = false;
    if (SMALL) {
        __syncthreads();                 // mbarrier initialised
        if (tid == 0) {
            const unsigned b_start = (unsigned)(p.n_frag_padded * 8);
            const unsigned b_sn = (unsigned)(p.n_frag_padded * 4);
            const unsigned b_grid = (unsigned)(p.grid_padded * 2);
            const unsigned b_desc = (unsigned)(p.n_tracks_padded * 16);
            const unsigned b_hint = (unsigned)(p.n_tracks_padded * 32);
            const unsigned b_base = (unsigned)(p.n_base_padded * 4);
            mbar_expect_tx(&sm.mbar, b_start + b_sn + b_grid + b_desc + b_hint + b_base);
            unsigned char *d = dyn_smem;
            bulk_g2s_chunked(d, p.start, b_start, &sm.mbar); d += b_start;
            bulk_g2s_chunked(d, p.sn, b_sn, &sm.mbar); d += b_sn;
            bulk_g2s_chunked(d, p.grid, b_grid, &sm.mbar); d += b_grid;
            bulk_g2s_chunked(d, p.desc, b_desc, &sm.mbar); d += b_desc;
            bulk_g2s_chunked(d, p.hint, b_hint, &sm.mbar); d += b_hint;
            bulk_g2s_chunked(d, p.playlist_base, b_base, &sm.mbar);
        }
        stage_pending = true;
    }

    // ---- tickets: thread 0 asks for the next tile right after publishing the current one's
    //      aggregate, so the atomic's L2 round trip hides behind the emission that follows.
    //      (No deeper look-ahead: a claimed tile is a tile no idle CTA can take.)
    unsigned t_new = 0;
    unsigned slot = 0;
    // Programmatic dependent launch: let the next batch's CTAs take the slots this grid frees
    // while it drains (only takes effect when the host launched with that attribute).
    asm volatile("griddepcontrol.launch_dependents;");
    Scan::acquire_set(p.ctr, p.tslot);
    if (tid == 0) sm.tile_slot[0] = (int)atomicAdd(&p.ctr->ticket[p.tslot & 1u], 1u);
    __syncthreads();
    // Tickets number the tiles of all the launch's batches consecutively: gB is the ticket, bB the batch
    // it falls in, tileB the tile's number inside that batch.  (Tickets grow, so bB only moves forward.)
    int gB = sm.tile_slot[0];
    int bB = 0;
    if constexpr (FUSED)
        while (bB + 1 < nb && gB >= p.fb[bB + 1].tile_first) bB++;
    int tileB = FUSED ? gB - p.fb[bB].tile_first : gB;
    int bA = 0;                      // batch of the tile waiting to be emitted

    // ---- the tile searched in the previous iteration, waiting to be emitted -----------------
    int tileA = -1;
    unsigned long long a_rel = 0, a_cnt_tile = 0;
    int a_cnt[kIpt];
    using SrcT = typename std::conditional<SMALL, int, long long>::type;
    SrcT a_src[kIpt];
    int a_snf[kIpt];                 // SNC: sn of each query's first kept fragment
    int a_rank = 0;
#pragma unroll
    for (int k = 0; k < kIpt; k++) { a_cnt[k] = 0; a_src[k] = 0; a_snf[k] = 0; }

    bool ticket_pending = false;         // thread 0: a ticket was requested for the next tile
    for (;;) {
        const bool haveB = gB < p.n_tiles;
        if (!haveB && tileA < 0) break;
        const BatchDesc &A = p.fb[FUSED ? bA : 0];           // (kernel parameters: constant bank, indexed)
        const BatchDesc &B = p.fb[FUSED ? bB : 0];
        unsigned long long *const statusA = FUSED ? p.tile_status + A.tile_first : p.tile_status;

        // tile A's look-back window: loaded between B's bounds and B's scan, used after the scan
        // (hides the L2 trip)
        unsigned long long pre[1] = {0ull};
        if (tileA > 0 && !haveB && warp == 0) pre[0] = ld_relaxed_u64(&statusA[max(tileA - 1 - lane, 0)]);

        // ---- tile A's look-back, by warp 0 alone, 32 predecessors per step, no CTA barrier.  (Run
        //      earlier, while tile B's window is landing in the general kernel, it measured the same.)
        auto lookback = [&]() {
            if (warp == 0) {
#if HLSMAP_TRACE_BUILD
                int diag = 0;
                const unsigned long long b = Scan::warp_prefix(statusA, p.epoch, tileA, a_cnt_tile, pre[0], A.base, &diag);
                if (p.trace != nullptr && lane == 0) p.trace[(long long)(A.tile_first + tileA) * 16 + 7] = (unsigned long long)diag;
#else
                const unsigned long long b = Scan::warp_prefix(statusA, p.epoch, tileA, a_cnt_tile, pre[0], A.base);
#endif
                if (lane == 0) sm.tile_base = b;
                if (SMALL) trace_mark_after(p, A.tile_first + tileA, 14, (int)b);
            }
        };
        unsigned long long tile_base = 0;
        const bool expand = tileA >= 0 && A.sn_out != nullptr && a_cnt_tile > 0 && a_cnt_tile <= 0x7fffffffull;
        constexpr int kRecShift = RecTable<SMALL, S::kTile>::kShift;
        const unsigned sn_addr = SMALL ? smem_u32(s_sn) : 0u;
        // start address of slot q = 2 * (sn address of slot q) + start_bias
        const unsigned start_bias = SMALL ? smem_u32(s_start) - 2u * sn_addr : 0u;
        // mark where this thread's records start inside the window [w0, w0 + kWindow)
        auto mark_heads = [&](uint2 *H, int w0) {
            int r2 = (int)a_rel;
#pragma unroll
            for (int k = 0; k < kIpt; k++) {
                if (a_cnt[k] > 0) {
                    const unsigned d = (unsigned)(r2 - w0);
                    if (d < (unsigned)S::kBits) atomicOr(&H[d >> 5].x, 1u << (d & 31));
                    r2 += a_cnt[k];
                }
            }
        };
        // ---- tile A, before its hits can be written: expansion records and the first window's head
        //      marks (no prefix needed: the other warps lay them down while warp 0 looks back), one
        //      barrier that hands out both, offsets[]
        auto emit_prep = [&]() {
            const long long q0 = (long long)tileA * p.tile_q + tid * kIpt;
            const long long q_end = min(A.nq, (long long)(tileA + 1) * p.tile_q);
            const bool full = A.vec_ok && q0 + kIpt <= q_end;
            trace_mark(p, A.tile_first + tileA, 3);
            lookback();
            if (expand) {   // records: one per query that kept >= 1 fragment, in query order
                int rank = a_rank, rel = (int)a_rel;
#pragma unroll
                for (int k = 0; k < kIpt; k++) {
                    if (a_cnt[k] > 0) {
                        sm.rec.set(rank, rel, a_src[k], sn_addr, a_snf[k]);
                        rank++;
                        rel += a_cnt[k];
                    }
                }
                mark_heads(sm.hb[hbuf], 0);
            }
            __syncthreads();
            tile_base = sm.tile_base;
            trace_mark(p, A.tile_first + tileA, 4);

            // ---- offsets[q] --------------------------------------------------------------
            {
                unsigned long long run = tile_base + a_rel;
                int o[kIpt];
#pragma unroll
                for (int k = 0; k < kIpt; k++) {
                    o[k] = (int)run;
                    run += (unsigned long long)a_cnt[k];
                }
                if (full && kIpt == 4) {
                    __stcs(reinterpret_cast<int4 *>(A.offsets + q0), make_int4(o[0], o[1], o[2], o[kIpt - 1]));
                } else if (full && kIpt == 2) {
                    __stcs(reinterpret_cast<int2 *>(A.offsets + q0), make_int2(o[0], o[1]));
                } else {
#pragma unroll
                    for (int k = 0; k < kIpt; k++)
                        if (q0 + k < q_end) A.offsets[q0 + k] = o[k];
                }
            }
            if (tileA == A.n_tiles - 1 && tid == kThreads - 1) {
                const unsigned long long total_hits = tile_base + a_cnt_tile;
                A.offsets[A.nq] = (int)total_hits;
                if (!FUSED || bA == nb - 1) {       // the context reports its most recent batch
                    p.result->total_hits = (long long)total_hits;
                    p.result->overflow = total_hits > 0x7fffffffull ? 1 : 0;
                }
            }
        };

        // =====================================================================================
        //  SEARCH tile B
        // =====================================================================================
        long long off[kIpt];
        int lo[kIpt], cnt[kIpt], snf[kIpt];
        unsigned long long b_rel = 0, b_cnt_tile = 0;
        int b_rank = 0;
#pragma unroll
        for (int k = 0; k < kIpt; k++) { off[k] = 0; lo[k] = 0; cnt[k] = 0; snf[k] = 0; }

        if (haveB) {
            // a tile holds p.tile_q <= 256*kIpt queries; threads past its end idle through the search
            const int tq0 = tid * kIpt;
            const long long q0 = (long long)tileB * p.tile_q + tq0;
            const long long q_end = min(B.nq, (long long)(tileB + 1) * p.tile_q);
            const bool full = B.vec_ok && q0 + kIpt <= q_end;
            trace_mark(p, gB, 0);
            {   // Pull the queries of the tile one grid-width ahead towards L2.  Tiles are handed
                // out in order, so every tile beyond the first wave is prefetched exactly once,
                // about one round before whichever CTA claims it loads them.  (Pulling that tile's
                // window of start[] towards L2 as well -- first / last playlist -> descriptors -> span,
                // followed by one warp over the iteration -- measured slower: 0.295 against 0.283 ms
                // per configs[4] chunk; predicating the probes of finished chains in the bounds
                // measured the same; staging tile A's window a second time, after B's bounds, so that
                // A's hits are gathered from shared memory instead of L2: 0.329 ms.)
                const long long gC = (long long)gB + gridDim.x;
                if (gC < p.n_tiles) {
                    int bC = bB;
                    if constexpr (FUSED)
                        while (bC + 1 < nb && gC >= p.fb[bC + 1].tile_first) bC++;
                    const BatchDesc &C = p.fb[FUSED ? bC : 0];
                    const long long c0 = (FUSED ? gC - C.tile_first : gC) * p.tile_q;
                    const long long cn = min((long long)p.tile_q, C.nq - c0);
                    const int l64 = (int)((cn * 8 + 127) >> 7), l32 = (int)((cn * 4 + 127) >> 7);
                    const char *a = nullptr;
                    if (tid < l64) a = reinterpret_cast<const char *>(C.q_begin + c0) + tid * 128;
                    else if (tid < 2 * l64) a = reinterpret_cast<const char *>(C.q_duration + c0) + (tid - l64) * 128;
                    else if (tid < 2 * l64 + l32) a = reinterpret_cast<const char *>(C.q_level + c0) + (tid - 2 * l64) * 128;
                    else if (tid < 2 * l64 + 2 * l32 && C.q_playlist)
                        a = reinterpret_cast<const char *>(C.q_playlist + c0) + (tid - 2 * l64 - l32) * 128;
                    if (a) asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
                }
            }

            // ---- load this thread's consecutive queries -----------------------------------
            int pl[kIpt], lv[kIpt];
            double bg[kIpt], du[kIpt];
            if (full) {
                if (kIpt == 4) {
                    const int4 l4 = __ldcs(reinterpret_cast<const int4 *>(B.q_level + q0));
                    lv[0] = l4.x; lv[1] = l4.y; lv[2] = l4.z; lv[kIpt - 1] = l4.w;
                    if (B.q_playlist) {
                        const int4 p4 = __ldcs(reinterpret_cast<const int4 *>(B.q_playlist + q0));
                        pl[0] = p4.x; pl[1] = p4.y; pl[2] = p4.z; pl[kIpt - 1] = p4.w;
                    }
                } else {
                    const int2 l2 = __ldcs(reinterpret_cast<const int2 *>(B.q_level + q0));
                    lv[0] = l2.x; lv[1] = l2.y;
                    if (B.q_playlist) {
                        const int2 p2 = __ldcs(reinterpret_cast<const int2 *>(B.q_playlist + q0));
                        pl[0] = p2.x; pl[1] = p2.y;
                    }
                }
                if (!B.q_playlist) {
#pragma unroll
                    for (int k = 0; k < kIpt; k++) pl[k] = 0;
                }
#pragma unroll
                for (int k = 0; k < kIpt; k += 2) {
                    const double2 b2 = __ldcs(reinterpret_cast<const double2 *>(B.q_begin + q0 + k));
                    const double2 d2 = __ldcs(reinterpret_cast<const double2 *>(B.q_duration + q0 + k));
                    bg[k] = b2.x; bg[k + 1] = b2.y;
                    du[k] = d2.x; du[k + 1] = d2.y;
                }
            } else {
#pragma unroll
                for (int k = 0; k < kIpt; k++) {
                    const bool in = q0 + k < q_end;
                    lv[k] = in ? __ldcs(B.q_level + q0 + k) : -1;
                    pl[k] = (in && B.q_playlist) ? __ldcs(B.q_playlist + q0 + k) : 0;
                    bg[k] = in ? __ldcs(B.q_begin + q0 + k) : 0.0;
                    du[k] = in ? __ldcs(B.q_duration + q0 + k) : 0.0;
                }
            }

            trace_mark_after(p, gB, 6, lv[0] ^ __double2loint(bg[kIpt - 1]) ^ __double2loint(du[kIpt - 1]));
            // ---- resolve (playlist, level) -> fragment table -----------------------------
            if (SMALL && stage_pending) {      // first tile only: tables + metadata have landed
                mbar_wait(&sm.mbar, mbar_phase);
                mbar_phase ^= 1u;
                stage_pending = false;
            }
            int len[kIpt], st[kIpt], trk[kIpt];
            int n_unp = 0, n_mis = 0;
            long long mn = 0x7fffffffffffffffll, mx = -1;
#pragma unroll
            for (int k = 0; k < kIpt; k++) {
                unsigned fl;
                st[k] = lookup_track<SMALL>(SMALL ? s_desc : p.desc, SMALL ? s_base : p.playlist_base,
                                            p.n_playlists, pl[k], lv[k], off[k], len[k], fl, &trk[k],
                                            p.uniform_levels);
                const bool in = q0 + k < q_end;
                n_unp += (in && st[k] == HLSMAP_ST_UNPARSED);
                n_mis += (in && st[k] == HLSMAP_ST_MISSING);
                if (!SMALL && st[k] == HLSMAP_ST_OK && len[k] > 0) {
                    mn = min(mn, off[k]);
                    mx = max(mx, off[k] + len[k]);
                }
            }
            if (!SMALL) trace_mark_after(p, gB, 13, len[0] ^ len[kIpt - 1]);
            if (B.status) {
                if (full && kIpt == 4) {
                    const unsigned w = (unsigned)st[0] | ((unsigned)st[1] << 8) |
                                       ((unsigned)st[2] << 16) | ((unsigned)st[kIpt - 1] << 24);
                    *reinterpret_cast<unsigned *>(B.status + q0) = w;
                } else {
#pragma unroll
                    for (int k = 0; k < kIpt; k++)
                        if (q0 + k < q_end) B.status[q0 + k] = (unsigned char)st[k];
                }
            }

            if (SMALL) {
                small_bounds<kIpt>(s_start, s_grid, s_hint, bg, du, off, len, st, trk, lo, cnt);
            } else {
                // ---- bounds: [lo, lo+cnt) is exactly what the filter keeps --------------------
                // The tile's tracks are searched from a shared-memory window of start[].  Usually
                // one window holds them all; a tile whose tracks span a few windows is done in
                // several passes (stage the window that begins at the first track still pending,
                // search every query whose track lies inside, repeat), so a query of such a tile
                // does not pay L2/HBM latency per probe.  Tracks longer than the window, and tiles
                // whose tracks lie far apart (queries not grouped by playlist), probe global memory.
                constexpr int kMaxPasses = 8;
                bool pend[kIpt];
                double ub[kIpt];
                const double *col[kIpt];
                int nn[kIpt], l[kIpt], h[kIpt];
#pragma unroll
                for (int k = 0; k < kIpt; k++) {
                    ub[k] = __dadd_rn(bg[k], du[k]);
                    pend[k] = st[k] == HLSMAP_ST_OK && len[k] > 0;
                }
                for (int pass = 0; pass < kMaxPasses; pass++) {
                    mn = 0x7fffffffffffffffll;
                    mx = -1;
#pragma unroll
                    for (int k = 0; k < kIpt; k++) {
                        if (pend[k] && len[k] <= p.stage_elems - 4) {
                            mn = min(mn, off[k]);
                            mx = max(mx, off[k] + len[k]);
                        }
                    }
#pragma unroll
                    for (int d = 16; d > 0; d >>= 1) {
                        mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, d));
                        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, d));
                    }
                    if (lane == 0) { sm.red_min[pass & 1][warp] = mn; sm.red_max[pass & 1][warp] = mx; }
                    __syncthreads();     // also: every thread is done with the previous window
#pragma unroll
                    for (int w = 0; w < kWarps; w++) {
                        mn = min(mn, sm.red_min[pass & 1][w]);
                        mx = max(mx, sm.red_max[pass & 1][w]);
                    }
                    if (mx < 0) break;                  // nothing (left) to stage: CTA-uniform
                    const long long need_lo = mn & ~3ll, need_all = (mx + 3) & ~3ll;
                    if (pass == 0 && need_all - need_lo > 4ll * p.stage_elems) break;   // far apart
                    const long long need_hi = min(need_all, need_lo + (long long)p.stage_elems);
                    if (!(need_lo >= span_lo && need_hi <= span_hi)) {   // else: still resident
                        span_lo = need_lo;
                        span_hi = need_hi;
                        if (tid == 0) {
                            const unsigned bytes = (unsigned)((span_hi - span_lo) * 8);
                            mbar_expect_tx(&sm.mbar, bytes);
                            bulk_g2s_chunked(dyn_smem, p.start + span_lo, bytes, &sm.mbar);
                        }
                        mbar_wait(&sm.mbar, mbar_phase);
                        mbar_phase ^= 1u;
                    }
                    if (pass == 0) trace_mark(p, gB, 14);
#pragma unroll
                    for (int k = 0; k < kIpt; k++) {
                        // (off >= span_lo: the window starts at the first pending track)
                        const bool act = pend[k] && len[k] <= p.stage_elems - 4 && off[k] + len[k] <= span_hi;
                        col[k] = s_start + (act ? (int)(off[k] - span_lo) : 0);
                        nn[k] = act ? len[k] : 0;
                    }
                    bounds_interleaved<kIpt, false>(col, nn, bg, ub, l, h);
#pragma unroll
                    for (int k = 0; k < kIpt; k++) {
                        if (nn[k] > 0) {
                            lo[k] = l[k];
                            cnt[k] = max(h[k] - l[k], 0);
                            pend[k] = false;
                        }
                    }
                    if (need_hi == need_all) break;     // the window held every pending track
                }
                // whatever is still pending probes L2/HBM (no lane has any in the common case)
#pragma unroll
                for (int k = 0; k < kIpt; k++) {
                    col[k] = p.start + (pend[k] ? off[k] : 0);
                    nn[k] = pend[k] ? len[k] : 0;
                }
                bounds_interleaved<kIpt, true>(col, nn, bg, ub, l, h);
#pragma unroll
                for (int k = 0; k < kIpt; k++) {
                    if (pend[k]) {
                        lo[k] = l[k];
                        cnt[k] = max(h[k] - l[k], 0);
                    }
                }
            }

            trace_mark_after(p, gB, 1, cnt[0] ^ cnt[kIpt - 1]);
            if (SNC) {       // consumed one iteration later, when this tile is emitted
#pragma unroll
                for (int k = 0; k < kIpt; k++) snf[k] = cnt[k] > 0 ? __ldg(p.sn + off[k] + lo[k]) : 0;
            }
            if (tileA > 0 && warp == 0) pre[0] = ld_relaxed_u64(&statusA[max(tileA - 1 - lane, 0)]);
            // ---- fused scan of (hit count, non-empty flag): thread -> warp -> CTA ---------
            unsigned long long tval = 0;
#pragma unroll
            for (int k = 0; k < kIpt; k++)
                tval += (unsigned long long)cnt[k] + ((unsigned long long)(cnt[k] > 0) << kCntBits);
            unsigned long long wtot;
            const unsigned long long texcl = warp_excl_scan64(tval, wtot);
            if (lane == 0) sm.warp_cnt[warp] = wtot;
            __syncthreads();     // also: everyone is past the previous emission of rec[]/hb[]
            unsigned long long wbase = 0, tile_tot = 0;
#pragma unroll
            for (int w = 0; w < kWarps; w++) {
                const unsigned long long c = sm.warp_cnt[w];
                wbase += (w < warp) ? c : 0ull;
                tile_tot += c;
            }
            trace_mark_after(p, gB, 8, (int)tile_tot);
            b_cnt_tile = tile_tot & kCntMask;
            Scan::publish_aggregate(FUSED ? p.tile_status + B.tile_first : p.tile_status, p.epoch, tileB, b_cnt_tile, B.base);
            trace_mark(p, gB, 2);
            b_rel = (wbase + texcl) & kCntMask;
            b_rank = (int)((wbase + texcl) >> kCntBits);
            {   // rare: fold the unparsed / missing tallies (one atomic per warp when non-zero)
                const unsigned any = __ballot_sync(0xffffffffu, (n_unp | n_mis) != 0 && (!FUSED || bB == nb - 1));
                if (any) {
#pragma unroll
                    for (int d = 16; d > 0; d >>= 1) {
                        n_unp += __shfl_xor_sync(0xffffffffu, n_unp, d);
                        n_mis += __shfl_xor_sync(0xffffffffu, n_mis, d);
                    }
                    if (lane == 0) {
                        if (n_unp) atomicAdd(&p.ctr->n_unparsed, (unsigned long long)n_unp);
                        if (n_mis) atomicAdd(&p.ctr->n_missing, (unsigned long long)n_mis);
                    }
                }
            }
        }

        // =====================================================================================
        //  EMIT tile A (searched one iteration ago): look-back, offsets, hits
        // =====================================================================================
        if (tileA >= 0) {
            emit_prep();

            // ---- expansion: the tile's hits form ONE contiguous output run; the whole CTA
            //      writes it 256 slots at a time whatever the per-query counts (load balanced)
            if (expand) {
                const int total = (int)a_cnt_tile;
                // shared-memory address of the record BEFORE the first one (ranks count from 1 below)
                const unsigned rec_addr = smem_u32(&sm.rec.v[0]) - (1u << kRecShift);
                unsigned sb_bias = 0;       // general kernel: address of sb[r] = ((address of v[r]) >> 1) + sb_bias
                if constexpr (!SMALL) sb_bias = smem_u32(&sm.rec.sb[0]) - (smem_u32(&sm.rec.v[0]) >> 1);
                int carried = 0;            // records that start before the current window
                const bool all_fit = (long long)tile_base + total <= A.capacity;   // CTA-uniform
                for (int w0 = 0; w0 < total; w0 += kWindow) {
                    uint2 *H = sm.hb[hbuf];
                    {   // head marks (the first window's are down already), rank words
                        if (w0 > 0) {
                            mark_heads(H, w0);
                            __syncthreads();
                        }
                        const int c0 = tid < kWords ? __popc(H[tid].x) : 0;
                        int wsum;
                        const int ex = warp_excl_scan(c0, wsum);
                        if (lane == 0) sm.warp_heads[warp] = wsum;
                        __syncthreads();
                        int before = carried;
#pragma unroll
                        for (int w = 0; w < kWarps; w++) before += (w < warp) ? sm.warp_heads[w] : 0;
                        // .y: where the record that covers the word's first slot lives (the records
                        // starting at or before a lane's slot are counted on top of it)
                        if (tid < kWords) H[tid].y = rec_addr + ((unsigned)(before + ex) << kRecShift);
                        // the next window starts kWindow slots on, inside this bitmap
                        if (tid == kWindow / 32) sm.carried = before + ex;
                        __syncthreads();
                        carried = sm.carried;
                    }
                    if (w0 == 0) trace_mark(p, A.tile_first + tileA, 9);
                    // The run is cut into chunks of kChunk slots that start at multiples of kChunk in the
                    // OUTPUT arrays (slot + a is a multiple of kChunk at a chunk's start): whatever the tile's
                    // first slot, every store instruction of a warp then covers whole 128-byte lines (two of
                    // time_out, one of sn_out) instead of straddling three and two -- measured 35.1 -> 33.0 us
                    // per batch of 2^20 queries.  Each warp takes every kWarps-th chunk of the window, its
                    // lanes 4 slots each, 32 apart.  The chunks cut by the run's two ends, and every chunk
                    // when the hits overflow the capacity, go slot by slot.
                    const int a = (int)(tile_base & (unsigned long long)(kChunk - 1));
                    // a lane's bit position inside its bitmap word is the same for all its slots
                    const unsigned below = (2u << ((unsigned)(lane - a) & 31u)) - 1u;
                    const int lim = min(w0 + kWindow, total);
                    const int k_lo = w0 == 0 ? 0 : (w0 + a + kChunk - 1) / kChunk;
                    const int k_hi = (lim + a + kChunk - 1) / kChunk;
                    // [f_lo, f_hi): the chunks that lie inside the run (and the capacity) with all their slots
                    int f_lo = k_lo + (k_lo * kChunk < a ? 1 : 0);
                    int f_hi = k_hi - (k_hi * kChunk - a > total ? 1 : 0);
                    if (!all_fit) f_hi = f_lo;
                    // (same in every thread; the reduction lets the compiler know, so the loops below
                    // are uniform and keep their memory descriptors in uniform registers)
                    f_hi = __reduce_max_sync(0xffffffffu, max(f_hi, f_lo));
                    auto slow_chunk = [&](int k) {
                        const int sl = k * kChunk - a + lane;
                        if constexpr (SMALL) {
#pragma unroll 1
                            for (int j = 0; j < 4; j++) {
                                const int q = sl + 32 * j;
                                if (q < 0 || q >= total || (long long)tile_base + q >= A.capacity) continue;
                                const uint2 h = H[(q - w0) >> 5];
                                const unsigned R = lds_u32<0>(h.y + 4u * (unsigned)__popc(h.x & below)) + 4u * (unsigned)q;
                                __stcs(A.time_out + ((long long)tile_base + q), lds_f64<0>(2u * R + start_bias));
                                __stcs(A.sn_out + ((long long)tile_base + q), (int)lds_u32<0>(R));
                            }
                        } else {        // all four loads first: they come from L2
                            double t[4];
                            int sv[4];
                            bool ok[4];
#pragma unroll
                            for (int j = 0; j < 4; j++) {
                                const int q = sl + 32 * j;
                                ok[j] = q >= 0 && q < total && (long long)tile_base + q < A.capacity;
                                t[j] = 0.0;
                                sv[j] = 0;
                                if (ok[j]) {
                                    const uint2 h = H[(q - w0) >> 5];
                                    const unsigned ra = h.y + 8u * (unsigned)__popc(h.x & below);
                                    const long long src = lds_s64<0>(ra) + q;
                                    t[j] = __ldg(p.start + src);
                                    sv[j] = SNC ? (int)lds_u32<0>((ra >> 1) + sb_bias) + q : __ldg(p.sn + src);
                                }
                            }
#pragma unroll
                            for (int j = 0; j < 4; j++) {
                                if (ok[j]) {
                                    __stcs(A.time_out + ((long long)tile_base + sl + 32 * j), t[j]);
                                    __stcs(A.sn_out + ((long long)tile_base + sl + 32 * j), sv[j]);
                                }
                            }
                        }
                    };
#pragma unroll 1
                    for (int k = k_lo + warp; k < f_lo; k += kWarps) slow_chunk(k);
                    // this lane's first slot of its warp's first full chunk, and what follows from it
                    int k = f_lo + warp;
                    const int sl0 = k * kChunk - a + lane;
                    const uint2 *hp = H + ((sl0 - w0) >> 5);
                    double *tp = A.time_out + ((long long)tile_base + sl0);
                    int *sp = A.sn_out + ((long long)tile_base + sl0);
                    unsigned slot4 = 4u * (unsigned)sl0;            // SMALL: 4 * slot
                    long long slotl = (long long)sl0;               // general: slot
                    auto body = [&]() {
                        double t[4];
                        int sv[4];
                        uint2 hw[4];
#pragma unroll
                        for (int j = 0; j < 4; j++) hw[j] = hp[j];
                        if constexpr (SMALL) {
                            unsigned R[4];
#pragma unroll
                            for (int j = 0; j < 4; j++)
                                R[j] = lds_u32<0>(hw[j].y + 4u * (unsigned)__popc(hw[j].x & below)) + slot4;
                            sv[0] = (int)lds_u32<0>(R[0]);
                            sv[1] = (int)lds_u32<128>(R[1]);
                            sv[2] = (int)lds_u32<256>(R[2]);
                            sv[3] = (int)lds_u32<384>(R[3]);
                            t[0] = lds_f64<0>(2u * R[0] + start_bias);
                            t[1] = lds_f64<256>(2u * R[1] + start_bias);
                            t[2] = lds_f64<512>(2u * R[2] + start_bias);
                            t[3] = lds_f64<768>(2u * R[3] + start_bias);
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; j++) {
                                const unsigned ra = hw[j].y + 8u * (unsigned)__popc(hw[j].x & below);
                                const long long src = lds_s64<0>(ra) + slotl + 32 * j;
                                t[j] = __ldg(p.start + src);      // L2-hot: the span was just bulk-copied
                                // sb[] lies behind v[]: entry r of v at rec_addr + 8 r  <=>  entry r of sb at sb_bias + (that >> 1)
                                sv[j] = SNC ? (int)lds_u32<0>((ra >> 1) + sb_bias) + (int)slotl + 32 * j : __ldg(p.sn + src);
                            }
                        }
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            __stcs(tp + 32 * j, t[j]);     // time = fragment.start (:48)
                            __stcs(sp + 32 * j, sv[j]);    // sn   = fragment.sn    (:46)
                        }
                        hp += kWarps * (kChunk / 32);
                        tp += kWarps * kChunk;
                        sp += kWarps * kChunk;
                        slot4 += 4u * kWarps * kChunk;
                        slotl += kWarps * kChunk;
                        k += kWarps;
                    };
                    // Claim the next tile as late as possible -- a claimed but unsearched tile stalls
                    // the look-back of every later tile -- yet early enough for the atomic's round
                    // trip to hide behind the slots still to write: kClaimSlots before the tile's last.
                    const bool claim_here = haveB && w0 + kWindow >= total;
                    const int k_claim = claim_here ? max(f_lo, f_hi - kClaimSlots / kChunk) : f_hi;
#pragma unroll 1
                    while (k < k_claim) body();
                    if (claim_here && tid == 0 && !ticket_pending) {
                        t_new = atomicAdd(&p.ctr->ticket[p.tslot & 1u], 1u);
                        ticket_pending = true;
                    }
#pragma unroll 1
                    while (k < f_hi) body();
                    if (w0 == 0) trace_mark(p, A.tile_first + tileA, 10);
#pragma unroll 1
                    for (k = f_hi + ((warp - (f_hi - k_lo)) & (kWarps - 1)); k < k_hi; k += kWarps) slow_chunk(k);
                    if (w0 == 0) trace_mark(p, A.tile_first + tileA, 11);
                    __syncthreads();
                    if (w0 == 0) trace_mark(p, A.tile_first + tileA, 12);
                    if (tid < kWords) H[tid].x = 0u;   // next used two windows from now
                    hbuf ^= 1u;
                }
            }
        }

        // ---- rotate the pipeline: B becomes the pending tile --------------------------------
        if (tileA >= 0) trace_mark(p, A.tile_first + tileA, 5);
        slot ^= 1u;            // alternate slots: the one read last iteration is not rewritten
        if (tid == 0) {
            if (haveB) {
                if (!ticket_pending) t_new = atomicAdd(&p.ctr->ticket[p.tslot & 1u], 1u);
                sm.tile_slot[slot] = (int)t_new;
                ticket_pending = false;
            } else {
                sm.tile_slot[slot] = p.n_tiles;
            }
        }
        __syncthreads();       // all global writes of tile A issued; next tile id visible
        tileA = haveB ? tileB : -1;
        if constexpr (FUSED) bA = bB;
        a_rel = b_rel;
        a_cnt_tile = b_cnt_tile;
        a_rank = b_rank;
#pragma unroll
        for (int k = 0; k < kIpt; k++) {
            a_cnt[k] = cnt[k];
            a_src[k] = (SrcT)(off[k] + lo[k]);
            a_snf[k] = snf[k];
        }
        gB = sm.tile_slot[slot];
        if constexpr (FUSED)
            while (bB + 1 < nb && gB >= p.fb[bB + 1].tile_first) bB++;
        tileB = FUSED ? gB - p.fb[bB].tile_first : gB;
    }
    // never leave with a bulk copy still landing in this CTA's shared memory
    if (stage_pending) mbar_wait(&sm.mbar, mbar_phase);
    // the last CTA to leave wraps the batch up (counters -> result, scratch header re-armed)
    Scan::leave_set(p.ctr, p.tslot, p.result);
}

// =========================================================================================
//  upload-time hint grid (SMALL maps): grid[c] = fragments of the track whose cell < c
// =========================================================================================
__global__ void __launch_bounds__(256)
build_hint_kernel(const double *__restrict__ start, const TrackDesc *__restrict__ desc,
                  TrackHint *hint, unsigned short *grid, int n_tracks)
{
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    for (int t = blockIdx.x * wpb + (threadIdx.x >> 5); t < n_tracks; t += gridDim.x * wpb) {
        const TrackDesc d = desc[t];
        TrackHint h = hint[t];            // grid_off and ncell were planned on the host
        const bool live = (d.flags & kTrackStateMask) == HLSMAP_TRACK_PARSED && d.len > 0 &&
                          (d.flags & kTrackMonotone);
        double t0 = 0.0, inv = 0.0;
        if (live) {
            t0 = start[d.off];
            const double span = __dsub_rn(start[d.off + d.len - 1], t0);
            inv = __ddiv_rn((double)h.ncell, span);
            if (!(span > 0.0) || !(inv > 0.0) || !(inv <= 1.7e308) || !(fabs(t0) <= 1.7e308))
                inv = 0.0;                // degenerate: every value lands in cell 0
        }
        if (lane == 0) {
            h.t0 = t0;
            h.inv = inv;
            hint[t] = h;
        }
        const GmemCol col{start + d.off};
        for (int c = lane; c <= h.ncell; c += 32) {
            int lo = 0, n = live ? d.len : 0;   // first i with cell(start[i]) >= c
            while (n > 0) {
                const int half = n >> 1;
                const bool keep_right = hint_cell(col(lo + half), t0, inv, h.ncell) < c;
                lo = keep_right ? lo + half + 1 : lo;
                n = keep_right ? n - half - 1 : half;
            }
            grid[h.grid_off + c] = (unsigned short)(c == h.ncell ? (live ? d.len : 0) : lo);
        }
    }
}

// =========================================================================================
//  filter kernel: the reference loop itself, warp per query
// =========================================================================================
namespace {
constexpr int kFThreads = 256;
constexpr int kFMaskWords = 32;
constexpr int kFWarps = kFThreads / 32;
using FScan = TileScan<kFThreads>;
static_assert(kFThreads == kFilterTileQueries, "filter tile = one query per thread");
}  // namespace

__global__ void __launch_bounds__(kFThreads)
resolve_filter_kernel(const ResolveParams p)
{
    __shared__ FScan::Smem scan;
    __shared__ unsigned long long warp_cnt[kFWarps];
    // which fragments pass 1 kept, 32 per word, for tracks of up to 32 * kFMaskWords fragments: pass 2
    // then visits only the words that hold a hit instead of comparing the whole track again
    __shared__ unsigned kept[kFWarps][32][kFMaskWords];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lane_lt = (1u << lane) - 1u;

    FScan::acquire_set(p.ctr, p.tslot);
    for (;;) {
        const int tile = FScan::next_tile(scan, p.ctr, p.tslot);
        if (tile >= p.n_tiles) break;
        const long long q = (long long)tile * kFThreads + tid;
        const bool in = q < p.nq;

        const int lvl = in ? p.q_level[q] : -1;
        const int pls = (in && p.q_playlist) ? p.q_playlist[q] : 0;
        const double bg = in ? p.q_begin[q] : 0.0;
        const double ub = in ? __dadd_rn(bg, p.q_duration[q]) : 0.0;
        long long off;
        int len;
        unsigned fl;
        const int st = lookup_track(p.desc, p.playlist_base, p.n_playlists, pls, lvl, off, len, fl);
        if (in && p.status) p.status[q] = (unsigned char)st;
        if (in && st == HLSMAP_ST_UNPARSED) atomicAdd(&p.ctr->n_unparsed, 1ull);
        if (in && st == HLSMAP_ST_MISSING) atomicAdd(&p.ctr->n_missing, 1ull);
        if (!in) len = 0;

        // pass 1: count -- the warp walks query j's fragments 32 at a time (:42-44)
        int mycnt = 0;
        for (int j = 0; j < 32; j++) {
            const int jlen = __shfl_sync(0xffffffffu, len, j);
            if (jlen == 0) continue;
            const long long joff = __shfl_sync(0xffffffffu, off, j);
            const double jbg = __shfl_sync(0xffffffffu, bg, j);
            const double jub = __shfl_sync(0xffffffffu, ub, j);
            int c = 0;
            for (int i0 = 0; i0 < jlen; i0 += 128) {       // four words of 32 fragments, loads in flight together
                double s[4];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int i = i0 + 32 * k + lane;
                    s[k] = i < jlen ? __ldg(p.start + joff + i) : __longlong_as_double(0x7ff8000000000000ll);
                }
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const bool keep = (jbg <= s[k]) && (s[k] <= jub);      // NaN past the end: never kept
                    const unsigned ball = __ballot_sync(0xffffffffu, keep);
                    const int w = (i0 >> 5) + k;
                    if (lane == 0 && w < kFMaskWords) kept[warp][j][w] = ball;
                    c += __popc(ball);
                }
            }
            if (lane == j) mycnt = c;
        }

        int wtotal;
        const int lexcl = warp_excl_scan(mycnt, wtotal);
        if (lane == 0) warp_cnt[warp] = (unsigned long long)wtotal;
        __syncthreads();
        unsigned long long wbase = 0, tile_total = 0;
#pragma unroll
        for (int w = 0; w < kFWarps; w++) {
            const unsigned long long c = warp_cnt[w];
            wbase += (w < warp) ? c : 0ull;
            tile_total += c;
        }
        const unsigned long long tile_base =
            FScan::exclusive_prefix(scan, p.tile_status, p.epoch, tile, tile_total, p.base);
        const unsigned long long my_base = tile_base + wbase + (unsigned long long)lexcl;
        if (in) p.offsets[q] = (int)my_base;
        if (tile == p.n_tiles - 1 && tid == kFThreads - 1) {
            const unsigned long long total = tile_base + tile_total;
            p.offsets[p.nq] = (int)total;
            p.result->total_hits = (long long)total;
            p.result->overflow = total > 0x7fffffffull ? 1 : 0;
        }

        // pass 2: ordered compaction with ballot/popc (:45-49)
        if (p.sn_out != nullptr && wtotal > 0) {
            for (int j = 0; j < 32; j++) {
                const int jcnt = __shfl_sync(0xffffffffu, mycnt, j);
                if (jcnt == 0) continue;
                const int jlen = __shfl_sync(0xffffffffu, len, j);
                const long long joff = __shfl_sync(0xffffffffu, off, j);
                const double jbg = __shfl_sync(0xffffffffu, bg, j);
                const double jub = __shfl_sync(0xffffffffu, ub, j);
                long long out = (long long)__shfl_sync(0xffffffffu, my_base, j);
                __syncwarp();
                for (int i0 = 0; i0 < jlen; i0 += 32) {
                    const int i = i0 + lane;
                    bool keep = false;
                    double s = 0.0;
                    unsigned ball;
                    if (i0 < 32 * kFMaskWords) {           // remembered from pass 1
                        ball = kept[warp][j][i0 >> 5];
                        if (ball == 0u) continue;          // warp-uniform: nothing kept in these 32
                        keep = (ball >> lane) & 1u;
                        if (keep) s = __ldg(p.start + joff + i);
                    } else {
                        if (i < jlen) {
                            s = __ldg(p.start + joff + i);
                            keep = (jbg <= s) && (s <= jub);
                        }
                        ball = __ballot_sync(0xffffffffu, keep);
                    }
                    if (keep) {
                        const long long o = out + __popc(ball & lane_lt);
                        if (o < p.capacity) {
                            p.time_out[o] = s;
                            p.sn_out[o] = __ldg(p.sn + joff + i);
                        }
                    }
                    out += __popc(ball);
                }
            }
        }

        if (FScan::tile_done(p.ctr, p.n_tiles)) FScan::finish_batch(p.ctr, p.result, p.tslot);
    }
    FScan::leave_set(p.ctr, p.tslot);
}

// =========================================================================================
//  range form: (first, count) per query, no scan
// =========================================================================================
__global__ void __launch_bounds__(256)
range_kernel(const RangeParams p)
{
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < p.nq;
         q += (long long)gridDim.x * blockDim.x) {
        const double bg = p.q_begin[q];
        const double ub = __dadd_rn(bg, p.q_duration[q]);
        long long off;
        int len;
        unsigned fl;
        int st = lookup_track(p.desc, p.playlist_base, p.n_playlists,
                              p.q_playlist ? p.q_playlist[q] : 0, p.q_level[q], off, len, fl);
        int first = -1, count = 0;
        if (st == HLSMAP_ST_OK) {
            if (fl & kTrackMonotone) {
                const GmemCol col{p.start + off};
                const int l = lower_bound_begin(col, 0, len, bg);
                const int h = upper_bound_from(col, l, len, ub);
                count = h - l;
                first = count > 0 ? l : -1;
            } else {
                int last = -1;
                for (int i = 0; i < len; i++) {
                    const double s = __ldg(p.start + off + i);
                    if (bg <= s && s <= ub) {
                        if (first < 0) first = i;
                        last = i;
                        count++;
                    }
                }
                if (count > 0 && last - first + 1 != count) st = HLSMAP_ST_SCATTERED;
            }
        }
        p.first[q] = first;
        p.count[q] = count;
        if (p.status) p.status[q] = (unsigned char)st;
    }
}

// Range form on a map that fits shared memory (SMALL): the same staging and hint-grid search as
// the list form's kernel, without its scan and expansion.  Every CTA stages start[], the hint grid
// and the track metadata once (bulk copies, sn[] is not needed), then walks 1024-query tiles
// grid-stride: four queries per thread, 128-bit loads and stores.
namespace {
constexpr int kRangeIpt = 4;
constexpr int kRangeTile = kThreads * kRangeIpt;
}  // namespace

__global__ void __launch_bounds__(kThreads)
range_small_kernel(const RangeParams p, const TrackHint *hint, const unsigned short *grid,
                   long long n_frag_padded, int grid_padded, int n_tracks_padded, int n_base_padded, int vec_ok)
{
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ unsigned long long mbar;
    const double *s_start = reinterpret_cast<const double *>(dyn_smem);
    const unsigned short *s_grid = reinterpret_cast<const unsigned short *>(dyn_smem + n_frag_padded * 8);
    const unsigned char *s_meta = dyn_smem + n_frag_padded * 8 + (long long)grid_padded * 2;
    const TrackDesc *s_desc = reinterpret_cast<const TrackDesc *>(s_meta);
    const TrackHint *s_hint = reinterpret_cast<const TrackHint *>(s_meta + (long long)n_tracks_padded * 16);
    const int *s_base = reinterpret_cast<const int *>(s_meta + (long long)n_tracks_padded * 48);
    const int tid = threadIdx.x;
    if (tid == 0) mbar_init(&mbar, 1);
    __syncthreads();
    if (tid == 0) {
        const unsigned b_start = (unsigned)(n_frag_padded * 8), b_grid = (unsigned)(grid_padded * 2);
        const unsigned b_desc = (unsigned)(n_tracks_padded * 16), b_hint = (unsigned)(n_tracks_padded * 32);
        const unsigned b_base = (unsigned)(n_base_padded * 4);
        mbar_expect_tx(&mbar, b_start + b_grid + b_desc + b_hint + b_base);
        unsigned char *d = dyn_smem;
        bulk_g2s_chunked(d, p.start, b_start, &mbar); d += b_start;
        bulk_g2s_chunked(d, grid, b_grid, &mbar); d += b_grid;
        bulk_g2s_chunked(d, p.desc, b_desc, &mbar); d += b_desc;
        bulk_g2s_chunked(d, hint, b_hint, &mbar); d += b_hint;
        bulk_g2s_chunked(d, p.playlist_base, b_base, &mbar);
    }
    bool staged = false;
    const long long n_tiles = (p.nq + kRangeTile - 1) / kRangeTile;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long q0 = tile * kRangeTile + (long long)tid * kRangeIpt;
        const bool full = vec_ok && q0 + kRangeIpt <= p.nq;
        int pl[kRangeIpt], lv[kRangeIpt];
        double bg[kRangeIpt], du[kRangeIpt];
        if (full) {
            const int4 l4 = __ldcs(reinterpret_cast<const int4 *>(p.q_level + q0));
            lv[0] = l4.x; lv[1] = l4.y; lv[2] = l4.z; lv[3] = l4.w;
            if (p.q_playlist) {
                const int4 p4 = __ldcs(reinterpret_cast<const int4 *>(p.q_playlist + q0));
                pl[0] = p4.x; pl[1] = p4.y; pl[2] = p4.z; pl[3] = p4.w;
            } else {
                pl[0] = pl[1] = pl[2] = pl[3] = 0;
            }
#pragma unroll
            for (int k = 0; k < kRangeIpt; k += 2) {
                const double2 b2 = __ldcs(reinterpret_cast<const double2 *>(p.q_begin + q0 + k));
                const double2 d2 = __ldcs(reinterpret_cast<const double2 *>(p.q_duration + q0 + k));
                bg[k] = b2.x; bg[k + 1] = b2.y;
                du[k] = d2.x; du[k + 1] = d2.y;
            }
        } else {
#pragma unroll
            for (int k = 0; k < kRangeIpt; k++) {
                const bool in = q0 + k < p.nq;
                lv[k] = in ? __ldcs(p.q_level + q0 + k) : -1;
                pl[k] = (in && p.q_playlist) ? __ldcs(p.q_playlist + q0 + k) : 0;
                bg[k] = in ? __ldcs(p.q_begin + q0 + k) : 0.0;
                du[k] = in ? __ldcs(p.q_duration + q0 + k) : 0.0;
            }
        }
        if (!staged) {                   // first tile only: the tables have landed
            mbar_wait(&mbar, 0);
            staged = true;
        }
        long long off[kRangeIpt];
        int len[kRangeIpt], st[kRangeIpt], trk[kRangeIpt], lo[kRangeIpt], cnt[kRangeIpt];
#pragma unroll
        for (int k = 0; k < kRangeIpt; k++) {
            unsigned fl;
            st[k] = lookup_track<true>(s_desc, s_base, p.n_playlists, pl[k], lv[k], off[k], len[k], fl, &trk[k]);
        }
        small_bounds<kRangeIpt>(s_start, s_grid, s_hint, bg, du, off, len, st, trk, lo, cnt);
        int first[kRangeIpt];
#pragma unroll
        for (int k = 0; k < kRangeIpt; k++) first[k] = cnt[k] > 0 ? lo[k] : -1;
        if (full) {
            __stcs(reinterpret_cast<int4 *>(p.first + q0), make_int4(first[0], first[1], first[2], first[3]));
            __stcs(reinterpret_cast<int4 *>(p.count + q0), make_int4(cnt[0], cnt[1], cnt[2], cnt[3]));
            if (p.status)
                *reinterpret_cast<unsigned *>(p.status + q0) =
                    (unsigned)st[0] | ((unsigned)st[1] << 8) | ((unsigned)st[2] << 16) | ((unsigned)st[3] << 24);
        } else {
#pragma unroll
            for (int k = 0; k < kRangeIpt; k++) {
                if (q0 + k < p.nq) {
                    p.first[q0 + k] = first[k];
                    p.count[q0 + k] = cnt[k];
                    if (p.status) p.status[q0 + k] = (unsigned char)st[k];
                }
            }
        }
    }
    if (!staged) mbar_wait(&mbar, 0);    // never leave with a bulk copy still landing
}

// =========================================================================================
//  upload-time monotonicity proof: one warp per track
// =========================================================================================
// Also proves, per track, that the sn column is contiguous: sn[i] == sn[0] + i in 32-bit wrap-around
// arithmetic (what hls.js numbers from #EXT-X-MEDIA-SEQUENCE always is); the search kernel then
// computes a hit's sn instead of reading it, bit for bit the value stored in the table.
__global__ void __launch_bounds__(256)
monotone_check_kernel(const double *__restrict__ start, const int *__restrict__ sn, TrackDesc *desc, int n_tracks)
{
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    for (int t = blockIdx.x * wpb + (threadIdx.x >> 5); t < n_tracks; t += gridDim.x * wpb) {
        TrackDesc d = desc[t];
        if ((d.flags & kTrackStateMask) != HLSMAP_TRACK_PARSED) continue;
        bool bad = false, gap = false;
        const unsigned sn0 = d.len > 0 ? (unsigned)sn[d.off] : 0u;
        for (int i = lane; i < d.len; i += 32) {
            const double s = start[d.off + i];
            bad |= (s != s);
            if (i > 0) bad |= (s < start[d.off + i - 1]);
            gap |= (unsigned)sn[d.off + i] != sn0 + (unsigned)i;
        }
        const unsigned any_bad = __ballot_sync(0xffffffffu, bad);
        const unsigned any_gap = __ballot_sync(0xffffffffu, gap);
        if (lane == 0) desc[t].flags = d.flags | (any_bad ? 0u : kTrackMonotone) | (any_gap ? 0u : kTrackSnContig);
    }
}

// =========================================================================================
//  live playlists: slide the fragment windows of many tracks in place (one warp per track).
//  The batch analogue of hls.js replacing level.details.fragments on a playlist reload, which
//  the reference then reads live (media-map.js:28,41).
// =========================================================================================
__global__ void __launch_bounds__(256)
slide_tracks_kernel(double *start, int *sn, double *dur0, TrackDesc *desc, long long n,
                    const int *track, const int *drop, const long long *append_off,
                    const double *new_start, const int *new_sn, const double *new_dur0)
{
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    for (long long e = (long long)blockIdx.x * wpb + (threadIdx.x >> 5); e < n;
         e += (long long)gridDim.x * wpb) {
        const int t = track[e];
        TrackDesc d = desc[t];
        const int k = min(max(drop[e], 0), d.len);          // fragments evicted at the front
        const int keep = d.len - k;
        const long long a0 = append_off[e];
        const int add = (int)(append_off[e + 1] - a0);
        // shift left by k: every chunk reads at or beyond everything written before it
        for (int i0 = 0; i0 < keep; i0 += 32) {
            const int i = i0 + lane;
            double s = 0.0;
            int q = 0;
            if (i < keep) {
                s = start[d.off + k + i];
                q = sn[d.off + k + i];
            }
            __syncwarp();
            if (i < keep && k > 0) {
                start[d.off + i] = s;
                sn[d.off + i] = q;
            }
            __syncwarp();
        }
        for (int j = lane; j < add; j += 32) {              // append the new fragments
            start[d.off + keep + j] = new_start[a0 + j];
            sn[d.off + keep + j] = new_sn[a0 + j];
        }
        __syncwarp();
        const int len = keep + add;
        // re-prove monotonicity of the whole column and the contiguity of its sn (the search kernel
        // relies on both flags)
        bool bad = false, gap = false;
        const unsigned sn0 = len > 0 ? (unsigned)sn[d.off] : 0u;
        for (int i = lane; i < len; i += 32) {
            const double s = start[d.off + i];
            bad |= (s != s);
            if (i > 0) bad |= (s < start[d.off + i - 1]);
            gap |= (unsigned)sn[d.off + i] != sn0 + (unsigned)i;
        }
        const unsigned any_bad = __ballot_sync(0xffffffffu, bad);
        const unsigned any_gap = __ballot_sync(0xffffffffu, gap);
        if (lane == 0) {
            unsigned fl = d.flags & ~(kTrackMonotone | kTrackSnContig);
            if ((fl & kTrackStateMask) == HLSMAP_TRACK_PARSED && !any_bad) fl |= kTrackMonotone;
            if ((fl & kTrackStateMask) == HLSMAP_TRACK_PARSED && !any_gap) fl |= kTrackSnContig;
            desc[t].len = len;
            desc[t].flags = fl;
            if (new_dur0) dur0[t] = new_dur0[e];
        }
    }
}

// =========================================================================================
//  getSegmentDuration (media-map.js:81-87): first fragment's duration of the level
// =========================================================================================
__global__ void __launch_bounds__(256)
segment_duration_kernel(const TrackDesc *desc, const int *playlist_base, int n_playlists,
                        const double *dur0, long long n, const int *q_playlist,
                        const int *q_level, double *out, unsigned char *status)
{
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n;
         q += (long long)gridDim.x * blockDim.x) {
        long long off;
        int len;
        unsigned fl;
        int track = 0;
        int st = lookup_track(desc, playlist_base, n_playlists, q_playlist ? q_playlist[q] : 0,
                              q_level[q], off, len, fl, &track);
        double v = __longlong_as_double(0x7ff8000000000000ll);
        if (st == HLSMAP_ST_OK) {
            if (len == 0) st = HLSMAP_ST_EMPTY;  // 'All segments should have a duration' (:86)
            else v = __ldg(dur0 + track);
        }
        out[q] = v;
        if (status) status[q] = (unsigned char)st;
    }
}

// =========================================================================================
//  view construction from fragments (p2p-loader-generator.js:145-147): per (playlist, level,
//  fragment index) -> SegmentView{sn: frag.sn, trackView{level, urlId}, time: frag.start}
// =========================================================================================
__global__ void __launch_bounds__(256)
views_from_fragments_kernel(const TrackDesc *desc, const int *playlist_base, int n_playlists,
                            const double *start, const int *sn, long long n, const int *q_playlist,
                            const int *q_level, const int *q_index, int *out_sn, double *out_time,
                            unsigned char *status)
{
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n;
         q += (long long)gridDim.x * blockDim.x) {
        long long off;
        int len;
        unsigned fl;
        int st = lookup_track(desc, playlist_base, n_playlists, q_playlist ? q_playlist[q] : 0,
                              q_level[q], off, len, fl);
        int s = 0;
        double t = __longlong_as_double(0x7ff8000000000000ll);
        if (st == HLSMAP_ST_OK) {
            const int i = q_index[q];
            if ((unsigned)i >= (unsigned)len) st = HLSMAP_ST_NO_FRAGMENT;   // fragments[i] undefined
            else {
                s = __ldg(sn + off + i);
                t = __ldg(start + off + i);
            }
        }
        out_sn[q] = s;
        out_time[q] = t;
        if (status) status[q] = (unsigned char)st;
    }
}

// =========================================================================================
//  getNextSegmentView -- EXTENSION, no counterpart in this reference (the method was removed
//  in 3.2.5, CHANGELOG.md:89).  Defined here as: the fragment that follows, in the same level,
//  the first fragment whose sn equals the given sn; found = 0 when there is none.
// =========================================================================================
__global__ void __launch_bounds__(256)
next_segment_view_kernel(const TrackDesc *desc, const int *playlist_base, int n_playlists,
                         const double *start, const int *sn, long long n, const int *q_playlist,
                         const int *q_level, const int *q_sn, int *out_sn, double *out_time,
                         unsigned char *found, unsigned char *status)
{
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n;
         q += (long long)gridDim.x * blockDim.x) {
        long long off;
        int len;
        unsigned fl;
        const int st = lookup_track(desc, playlist_base, n_playlists, q_playlist ? q_playlist[q] : 0,
                                    q_level[q], off, len, fl);
        int s = 0, idx = -1;
        double t = __longlong_as_double(0x7ff8000000000000ll);
        if (st == HLSMAP_ST_OK && len > 0) {
            const int want = q_sn[q];
            // contiguous numbering (the HLS norm) puts it at want - sn[0]; verified, not assumed
            const long long guess = (long long)want - (long long)__ldg(sn + off);
            if (guess >= 0 && guess < len && __ldg(sn + off + guess) == want) idx = (int)guess;
            for (int i = 0; i < len && (idx < 0 || i < idx); i++)      // first match wins
                if (__ldg(sn + off + i) == want) { idx = i; break; }
        }
        const bool ok = idx >= 0 && idx + 1 < len;
        if (ok) {
            s = __ldg(sn + off + idx + 1);
            t = __ldg(start + off + idx + 1);
        }
        out_sn[q] = s;
        out_time[q] = t;
        found[q] = ok ? 1 : 0;
        if (status) status[q] = (unsigned char)st;
    }
}

// =========================================================================================
//  launchers
// =========================================================================================
static int g_sm_count = 0;
static int sm_count()
{
    if (!g_sm_count) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (g_sm_count <= 0) g_sm_count = 148;
    }
    return g_sm_count;
}

template <bool SMALL, int IPT, bool SNC>
static int max_grid_of(size_t smem)
{
    int per_sm = 0;
    cudaFuncSetAttribute(resolve_search_kernel<SMALL, IPT, SNC, false>,
                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(resolve_search_kernel<SMALL, IPT, SNC, true>,
                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resolve_search_kernel<SMALL, IPT, SNC, true>,
                                                      kThreads, smem) != cudaSuccess || per_sm < 1)
        per_sm = 1;
    return per_sm * sm_count();
}

// (both sn variants of the general kernel get their shared-memory opt-in here; their occupancy is the same)
int resolve_search_max_grid(size_t smem, bool small, int ipt)
{
    if (small) return ipt == 2 ? max_grid_of<true, 2, false>(smem) : max_grid_of<true, 4, false>(smem);
    if (ipt == 2) return min(max_grid_of<false, 2, false>(smem), max_grid_of<false, 2, true>(smem));
    return min(max_grid_of<false, 4, false>(smem), max_grid_of<false, 4, true>(smem));
}

int resolve_search_tile(bool small, int ipt) { (void)small; return kThreads * (ipt == 2 ? 2 : 4); }

int resolve_filter_max_grid()
{
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resolve_filter_kernel, kFThreads,
                                                      0) != cudaSuccess || per_sm < 1)
        per_sm = 1;
    return per_sm * sm_count();
}

template <bool SMALL, int IPT, bool SNC, bool FUSED>
static cudaError_t launch_search_variant(const SearchParams<FUSED ? kMaxFused : 1> &p, int grid, size_t smem,
                                         cudaStream_t s, bool overlap_next)
{
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = overlap_next ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, resolve_search_kernel<SMALL, IPT, SNC, FUSED>, p);
}

template <bool FUSED>
static cudaError_t launch_search_any(const SearchParams<FUSED ? kMaxFused : 1> &p, int grid, size_t smem,
                                     cudaStream_t s, bool overlap_next)
{
    cudaError_t e;
    if (p.small && p.ipt == 2) e = launch_search_variant<true, 2, false, FUSED>(p, grid, smem, s, overlap_next);
    else if (p.small) e = launch_search_variant<true, 4, false, FUSED>(p, grid, smem, s, overlap_next);
    else if (p.ipt == 2 && p.sn_contig) e = launch_search_variant<false, 2, true, FUSED>(p, grid, smem, s, overlap_next);
    else if (p.ipt == 2) e = launch_search_variant<false, 2, false, FUSED>(p, grid, smem, s, overlap_next);
    else if (p.sn_contig) e = launch_search_variant<false, 4, true, FUSED>(p, grid, smem, s, overlap_next);
    else e = launch_search_variant<false, 4, false, FUSED>(p, grid, smem, s, overlap_next);
    return e != cudaSuccess ? e : cudaGetLastError();
}

// One batch, described by the launch-wide fields of p.
// overlap_next: this batch may begin while the previous kernel on the stream is draining.
cudaError_t launch_resolve_search(const ResolveParams &p, int grid, size_t smem, cudaStream_t s,
                                  bool overlap_next)
{
    SearchParams<1> sp{};
    static_cast<ResolveParams &>(sp) = p;
    sp.n_batches = 1;
    BatchDesc &b = sp.fb[0];
    b.q_playlist = p.q_playlist;
    b.q_level = p.q_level;
    b.q_begin = p.q_begin;
    b.q_duration = p.q_duration;
    b.offsets = p.offsets;
    b.sn_out = p.sn_out;
    b.time_out = p.time_out;
    b.status = p.status;
    b.nq = p.nq;
    b.capacity = p.capacity;
    b.base = p.base;
    b.tile_first = 0;
    b.n_tiles = p.n_tiles;
    b.vec_ok = p.vec_ok;
    return launch_search_any<false>(sp, grid, smem, s, overlap_next);
}

// n <= kMaxFused batches with one launch: p carries the tables, the scratch and the launch's total
// number of tiles; the batches' own fields are in b[].
cudaError_t launch_resolve_search_fused(const ResolveParams &p, const BatchDesc *b, int n, int grid, size_t smem,
                                        cudaStream_t s, bool overlap_next)
{
    SearchParams<kMaxFused> sp{};
    static_cast<ResolveParams &>(sp) = p;
    sp.n_batches = n;
    for (int i = 0; i < n; i++) sp.fb[i] = b[i];
    return launch_search_any<true>(sp, grid, smem, s, overlap_next);
}

cudaError_t launch_resolve_filter(const ResolveParams &p, int grid, cudaStream_t s)
{
    resolve_filter_kernel<<<grid, kFThreads, 0, s>>>(p);
    return cudaGetLastError();
}

// SMALL maps: the staged kernel.  smem = start + grid + descriptors + hints + playlist index.
cudaError_t launch_range_small(const RangeParams &p, const TrackHint *hint, const unsigned short *grid,
                               long long n_frag_padded, int grid_padded, int n_tracks_padded,
                               int n_base_padded, int vec_ok, cudaStream_t s)
{
    const size_t smem = (size_t)(n_frag_padded * 8 + (long long)grid_padded * 2 +
                                 (long long)n_tracks_padded * 48 + (long long)n_base_padded * 4);
    // the shared-memory opt-in and the occupancy belong to the current device: remembered per device
    static size_t configured[64] = {};
    static int per_sm_of[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    const int slot = dev & 63;
    if (smem != configured[slot]) {
        cudaError_t e = cudaFuncSetAttribute(range_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        int per_sm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, range_small_kernel, kThreads, smem) != cudaSuccess ||
            per_sm < 1)
            per_sm = 1;
        configured[slot] = smem;
        per_sm_of[slot] = per_sm;
    }
    long long blocks = (p.nq + kRangeTile - 1) / kRangeTile;
    int n_sm = 0;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    const long long cap = (long long)(n_sm > 0 ? n_sm : 148) * per_sm_of[slot];
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    range_small_kernel<<<(int)blocks, kThreads, smem, s>>>(p, hint, grid, n_frag_padded, grid_padded,
                                                            n_tracks_padded, n_base_padded, vec_ok);
    return cudaGetLastError();
}

cudaError_t launch_range(const RangeParams &p, cudaStream_t s)
{
    long long blocks = (p.nq + 255) / 256;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    range_kernel<<<(int)blocks, 256, 0, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_monotone_check(const double *start, const int *sn, TrackDesc *desc, int n_tracks,
                                  cudaStream_t s)
{
    if (n_tracks <= 0) return cudaSuccess;
    long long blocks = ((long long)n_tracks + 7) / 8;
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    monotone_check_kernel<<<(int)blocks, 256, 0, s>>>(start, sn, desc, n_tracks);
    return cudaGetLastError();
}

cudaError_t launch_build_hint(const double *start, const TrackDesc *desc, TrackHint *hint,
                              unsigned short *grid, int n_tracks, cudaStream_t s)
{
    if (n_tracks <= 0) return cudaSuccess;
    long long blocks = ((long long)n_tracks + 7) / 8;
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    build_hint_kernel<<<(int)blocks, 256, 0, s>>>(start, desc, hint, grid, n_tracks);
    return cudaGetLastError();
}

cudaError_t launch_slide_tracks(double *start, int *sn, double *dur0, TrackDesc *desc, long long n,
                                const int *track, const int *drop, const long long *append_off,
                                const double *new_start, const int *new_sn, const double *new_dur0,
                                cudaStream_t s)
{
    if (n <= 0) return cudaSuccess;
    long long blocks = (n + 7) / 8;
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    slide_tracks_kernel<<<(int)blocks, 256, 0, s>>>(start, sn, dur0, desc, n, track, drop, append_off,
                                                    new_start, new_sn, new_dur0);
    return cudaGetLastError();
}

cudaError_t launch_segment_duration(const TrackDesc *desc, const int *playlist_base,
                                    int n_playlists, const double *dur0, long long n,
                                    const int *q_playlist, const int *q_level, double *out,
                                    unsigned char *status, cudaStream_t s)
{
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    segment_duration_kernel<<<(int)blocks, 256, 0, s>>>(desc, playlist_base, n_playlists, dur0, n,
                                                        q_playlist, q_level, out, status);
    return cudaGetLastError();
}

cudaError_t launch_views_from_fragments(const TrackDesc *desc, const int *playlist_base, int n_playlists,
                                        const double *start, const int *sn, long long n,
                                        const int *q_playlist, const int *q_level, const int *q_index,
                                        int *out_sn, double *out_time, unsigned char *status,
                                        cudaStream_t s)
{
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    views_from_fragments_kernel<<<(int)blocks, 256, 0, s>>>(desc, playlist_base, n_playlists, start, sn, n,
                                                            q_playlist, q_level, q_index, out_sn,
                                                            out_time, status);
    return cudaGetLastError();
}

cudaError_t launch_next_segment_view(const TrackDesc *desc, const int *playlist_base, int n_playlists,
                                     const double *start, const int *sn, long long n,
                                     const int *q_playlist, const int *q_level, const int *q_sn,
                                     int *out_sn, double *out_time, unsigned char *found,
                                     unsigned char *status, cudaStream_t s)
{
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    next_segment_view_kernel<<<(int)blocks, 256, 0, s>>>(desc, playlist_base, n_playlists, start, sn, n,
                                                         q_playlist, q_level, q_sn, out_sn, out_time,
                                                         found, status);
    return cudaGetLastError();
}

}  // namespace hlsmap

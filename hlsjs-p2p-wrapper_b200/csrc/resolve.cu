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
constexpr int kIpt = 4;                         // queries per thread
constexpr int kTile = kThreads * kIpt;          // == kSearchTileQueries
constexpr int kWindow = 16384;                  // outputs expanded per pass over the tile
constexpr int kWords = kWindow / 32;            // head-bitmap words per window (2 per thread)
constexpr int kCntBits = 44;                    // fused scan: [nz:20][count:44]
constexpr unsigned long long kCntMask = (1ull << kCntBits) - 1ull;
static_assert(kTile == kSearchTileQueries, "tile size mismatch");
static_assert(kWords == 2 * kThreads, "two bitmap words per thread");

using Scan = TileScan<kThreads>;

// Expansion records: one per query that kept >= 1 fragment, in query order.
//   rel = first output slot relative to the tile's output run, src = first kept fragment.
template <bool SMALL>
struct RecTable;
template <>
struct RecTable<true> {   // every fragment index fits an int: one 8-byte record
    int2 v[kTile];
    __device__ __forceinline__ void set(int i, int rel, long long src) { v[i] = make_int2(rel, (int)src); }
    __device__ __forceinline__ void get(int i, int &rel, long long &src) const
    {
        const int2 r = v[i];
        rel = r.x;
        src = r.y;
    }
};
template <>
struct RecTable<false> {
    int rel_[kTile];
    long long src_[kTile];
    __device__ __forceinline__ void set(int i, int rel, long long src) { rel_[i] = rel; src_[i] = src; }
    __device__ __forceinline__ void get(int i, int &rel, long long &src) const
    {
        rel = rel_[i];
        src = src_[i];
    }
};

// Shared state of one CTA.
template <bool SMALL>
struct SearchSmem {
    Scan::Smem scan;
    unsigned long long mbar;
    long long red_min[kWarps];
    long long red_max[kWarps];
    unsigned long long warp_cnt[kWarps];
    int warp_heads[kWarps];
    RecTable<SMALL> rec;
    // per 32 output slots of the window: .x = bitmap of slots where a record starts,
    // .y = records that start before this word (tile-wide rank base).  Double buffered.
    uint2 hb[2][kWords];
};

__device__ __forceinline__ unsigned smem_u32(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
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
__device__ __forceinline__ int lookup_track(const TrackDesc *desc, const int *playlist_base,
                                            int n_playlists, int playlist, int level,
                                            long long &off, int &len, unsigned &flags,
                                            int *track = nullptr)
{
    off = 0;
    len = 0;
    flags = 0;
    if (track) *track = 0;
    if ((unsigned)playlist >= (unsigned)n_playlists) return HLSMAP_ST_MISSING;
    const int b = __ldg(playlist_base + playlist);
    const int e = __ldg(playlist_base + playlist + 1);
    if ((unsigned)level >= (unsigned)(e - b)) return HLSMAP_ST_MISSING;  // levels[level] undefined
    const int4 raw = __ldg(reinterpret_cast<const int4 *>(desc + b + level));
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

// =========================================================================================
//  search kernel.  SMALL = the whole map (start, sn, hint grid) lives in shared memory.
// =========================================================================================
template <bool SMALL>
__global__ void __launch_bounds__(kThreads)
resolve_search_kernel(const ResolveParams p)
{
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    __shared__ SearchSmem<SMALL> sm;
    double *s_start = reinterpret_cast<double *>(dyn_smem);
    const int *s_sn = reinterpret_cast<const int *>(dyn_smem + p.n_frag_padded * 8);
    const unsigned short *s_grid =
        reinterpret_cast<const unsigned short *>(dyn_smem + p.n_frag_padded * 12);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lane_le = 0xffffffffu >> (31 - lane);

    if (tid == 0) mbar_init(&sm.mbar, 1);
#pragma unroll
    for (int b = 0; b < 2; b++) {
        sm.hb[b][tid] = make_uint2(0u, 0u);
        sm.hb[b][tid + kThreads] = make_uint2(0u, 0u);
    }
    unsigned mbar_phase = 0;
    unsigned hbuf = 0;                   // which heads[] buffer the next window uses
    long long span_lo = 0, span_hi = 0;  // staged element range of start[]
    __syncthreads();

    bool stage_pending = false;
    if (SMALL) {
        span_hi = p.n_frag_padded;
        if (tid == 0) {
            const unsigned b_start = (unsigned)(p.n_frag_padded * 8);
            const unsigned b_sn = (unsigned)(p.n_frag_padded * 4);
            const unsigned b_grid = (unsigned)(p.grid_padded * 2);
            mbar_expect_tx(&sm.mbar, b_start + b_sn + b_grid);
            bulk_g2s_chunked(dyn_smem, p.start, b_start, &sm.mbar);
            bulk_g2s_chunked(dyn_smem + b_start, p.sn, b_sn, &sm.mbar);
            bulk_g2s_chunked(dyn_smem + b_start + b_sn, p.grid, b_grid, &sm.mbar);
        }
        stage_pending = true;
    }

    for (;;) {
        const int tile = Scan::next_tile(sm.scan, p.ctr, p.epoch);
        if (tile >= p.n_tiles) break;
        const long long q0 = (long long)tile * kTile + (long long)tid * kIpt;

        // ---- load this thread's 4 consecutive queries ------------------------------------
        int pl[kIpt], lv[kIpt];
        double bg[kIpt], du[kIpt];
        if (p.vec_ok && q0 + kIpt <= p.nq) {
            const int4 l4 = __ldcs(reinterpret_cast<const int4 *>(p.q_level + q0));
            lv[0] = l4.x; lv[1] = l4.y; lv[2] = l4.z; lv[3] = l4.w;
            if (p.q_playlist) {
                const int4 p4 = __ldcs(reinterpret_cast<const int4 *>(p.q_playlist + q0));
                pl[0] = p4.x; pl[1] = p4.y; pl[2] = p4.z; pl[3] = p4.w;
            } else {
                pl[0] = pl[1] = pl[2] = pl[3] = 0;
            }
            const double2 b0 = __ldcs(reinterpret_cast<const double2 *>(p.q_begin + q0));
            const double2 b1 = __ldcs(reinterpret_cast<const double2 *>(p.q_begin + q0 + 2));
            const double2 d0 = __ldcs(reinterpret_cast<const double2 *>(p.q_duration + q0));
            const double2 d1 = __ldcs(reinterpret_cast<const double2 *>(p.q_duration + q0 + 2));
            bg[0] = b0.x; bg[1] = b0.y; bg[2] = b1.x; bg[3] = b1.y;
            du[0] = d0.x; du[1] = d0.y; du[2] = d1.x; du[3] = d1.y;
        } else {
#pragma unroll
            for (int k = 0; k < kIpt; k++) {
                const bool in = q0 + k < p.nq;
                lv[k] = in ? __ldcs(p.q_level + q0 + k) : -1;
                pl[k] = (in && p.q_playlist) ? __ldcs(p.q_playlist + q0 + k) : 0;
                bg[k] = in ? __ldcs(p.q_begin + q0 + k) : 0.0;
                du[k] = in ? __ldcs(p.q_duration + q0 + k) : 0.0;
            }
        }

        // ---- resolve (playlist, level) -> fragment table --------------------------------
        long long off[kIpt];
        int len[kIpt], st[kIpt], trk[kIpt];
        int n_unp = 0, n_mis = 0;
        long long mn = 0x7fffffffffffffffll, mx = -1;
#pragma unroll
        for (int k = 0; k < kIpt; k++) {
            unsigned fl;
            st[k] = lookup_track(p.desc, p.playlist_base, p.n_playlists, pl[k], lv[k], off[k],
                                 len[k], fl, &trk[k]);
            const bool in = q0 + k < p.nq;
            n_unp += (in && st[k] == HLSMAP_ST_UNPARSED);
            n_mis += (in && st[k] == HLSMAP_ST_MISSING);
            if (!SMALL && st[k] == HLSMAP_ST_OK && len[k] > 0) {
                mn = min(mn, off[k]);
                mx = max(mx, off[k] + len[k]);
            }
        }
        if (p.status) {
            if (p.vec_ok && q0 + kIpt <= p.nq) {
                const unsigned w = (unsigned)st[0] | ((unsigned)st[1] << 8) |
                                   ((unsigned)st[2] << 16) | ((unsigned)st[3] << 24);
                *reinterpret_cast<unsigned *>(p.status + q0) = w;
            } else {
#pragma unroll
                for (int k = 0; k < kIpt; k++)
                    if (q0 + k < p.nq) p.status[q0 + k] = (unsigned char)st[k];
            }
        }

        // ---- make the needed span of start[] resident in shared memory -------------------
        bool staged = true;
        if (!SMALL) {
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, d));
                mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, d));
            }
            if (lane == 0) { sm.red_min[warp] = mn; sm.red_max[warp] = mx; }
            __syncthreads();
#pragma unroll
            for (int w = 0; w < kWarps; w++) {
                mn = min(mn, sm.red_min[w]);
                mx = max(mx, sm.red_max[w]);
            }
            const long long need_lo = mn & ~3ll, need_hi = (mx + 3) & ~3ll;
            if (mx < 0) {
                staged = true;  // nothing to search in this tile
            } else if (need_lo >= span_lo && need_hi <= span_hi) {
                staged = true;  // still resident from the previous tile
            } else if (need_hi - need_lo <= (long long)p.stage_elems) {
                span_lo = need_lo;
                span_hi = need_hi;
                if (tid == 0) {
                    const unsigned bytes = (unsigned)((span_hi - span_lo) * 8);
                    mbar_expect_tx(&sm.mbar, bytes);
                    bulk_g2s_chunked(dyn_smem, p.start + span_lo, bytes, &sm.mbar);
                }
                stage_pending = true;
            } else {
                staged = false;  // span larger than shared memory: search in L2/HBM
            }
        }
        if (stage_pending) {
            mbar_wait(&sm.mbar, mbar_phase);
            mbar_phase ^= 1u;
            stage_pending = false;
        }

        // ---- bounds: [lo, lo+cnt) is exactly what the filter keeps -----------------------
        int lo[kIpt], cnt[kIpt];
#pragma unroll
        for (int k = 0; k < kIpt; k++) {
            // one fp64 add, round-to-nearest, never contracted: JS `beginTime + duration`
            const double ub = __dadd_rn(bg[k], du[k]);
            int l, h;
            if (SMALL) {
                // hint grid narrows each bound to one cell; the filter's comparisons finish it
                const int4 h0 = __ldg(reinterpret_cast<const int4 *>(p.hint + trk[k]));
                const int4 h1 = __ldg(reinterpret_cast<const int4 *>(p.hint + trk[k]) + 1);
                const double t0 = __hiloint2double(h0.y, h0.x);
                const double inv = __hiloint2double(h0.w, h0.z);
                const int ncell = st[k] == HLSMAP_ST_OK ? h1.y : 1;
                const unsigned short *g = s_grid + (st[k] == HLSMAP_ST_OK ? h1.x : 0);
                const SmemCol col{s_start + (int)off[k]};
                const int cb = hint_cell(bg[k], t0, inv, ncell);
                const int cu = hint_cell(ub, t0, inv, ncell);
                const int b0 = g[cb], b1 = g[cb + 1];
                const int u0 = g[cu], u1 = g[cu + 1];
                l = lower_bound_begin(col, b0, b1 - b0, bg[k]);
                if (bg[k] != bg[k]) l = len[k];           // NaN: `begin <= start` never holds
                h = upper_bound_ub(col, u0, u1 - u0, ub);
                if (st[k] != HLSMAP_ST_OK) { l = 0; h = 0; }
            } else if (staged) {
                const SmemCol col{s_start + (int)(off[k] - span_lo)};
                l = lower_bound_begin(col, 0, len[k], bg[k]);
                h = upper_bound_from(col, l, len[k], ub);
            } else {
                const GmemCol col{p.start + off[k]};
                l = lower_bound_begin(col, 0, len[k], bg[k]);
                h = upper_bound_from(col, l, len[k], ub);
            }
            lo[k] = l;
            cnt[k] = max(h - l, 0);
        }

        // ---- fused scan of (hit count, non-empty flag): thread -> warp -> CTA -> batch ---
        unsigned long long tval = 0;
#pragma unroll
        for (int k = 0; k < kIpt; k++)
            tval += (unsigned long long)cnt[k] + ((unsigned long long)(cnt[k] > 0) << kCntBits);
        unsigned long long wtot;
        const unsigned long long texcl = warp_excl_scan64(tval, wtot);
        if (lane == 0) sm.warp_cnt[warp] = wtot;
        __syncthreads();
        unsigned long long wbase = 0, tile_tot = 0;
#pragma unroll
        for (int w = 0; w < kWarps; w++) {
            const unsigned long long c = sm.warp_cnt[w];
            wbase += (w < warp) ? c : 0ull;
            tile_tot += c;
        }
        {   // rare: fold the unparsed / missing tallies (one atomic per warp when non-zero)
            const unsigned any = __ballot_sync(0xffffffffu, (n_unp | n_mis) != 0);
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
        const unsigned long long tile_cnt = tile_tot & kCntMask;
        const unsigned long long tile_base =
            Scan::exclusive_prefix(sm.scan, p.tile_status, p.epoch, tile, tile_cnt);

        // ---- offsets[q] ------------------------------------------------------------------
        const unsigned long long my_rel = (wbase + texcl) & kCntMask;   // tile-relative
        {
            const unsigned long long my_base = tile_base + my_rel;
            int4 o4;
            o4.x = (int)my_base;
            o4.y = (int)(my_base + cnt[0]);
            o4.z = (int)(my_base + cnt[0] + cnt[1]);
            o4.w = (int)(my_base + cnt[0] + cnt[1] + cnt[2]);
            if (p.vec_ok && q0 + kIpt <= p.nq) {
                __stcs(reinterpret_cast<int4 *>(p.offsets + q0), o4);
            } else {
                if (q0 + 0 < p.nq) p.offsets[q0 + 0] = o4.x;
                if (q0 + 1 < p.nq) p.offsets[q0 + 1] = o4.y;
                if (q0 + 2 < p.nq) p.offsets[q0 + 2] = o4.z;
                if (q0 + 3 < p.nq) p.offsets[q0 + 3] = o4.w;
            }
        }
        if (tile == p.n_tiles - 1 && tid == kThreads - 1) {
            const unsigned long long total = tile_base + tile_cnt;
            p.offsets[p.nq] = (int)total;
            p.result->total_hits = (long long)total;
            p.result->overflow = total > 0x7fffffffull ? 1 : 0;
        }

        // ---- expansion: the tile's hits are ONE contiguous output run, written by the whole
        //      CTA 256 slots at a time whatever the per-query counts (load balanced) --------
        if (p.sn_out != nullptr && tile_cnt > 0 && tile_cnt <= 0x7fffffffull) {
            int rank = (int)((wbase + texcl) >> kCntBits);
            int rel = (int)my_rel;
#pragma unroll
            for (int k = 0; k < kIpt; k++) {
                if (cnt[k] > 0) {
                    sm.rec.set(rank, rel, off[k] + lo[k]);
                    rank++;
                    rel += cnt[k];
                }
            }
            const int total = (int)tile_cnt;
            int carried = 0;            // records that start before the current window
            for (int w0 = 0; w0 < total; w0 += kWindow) {
                uint2 *H = sm.hb[hbuf];
                {   // mark where this thread's records start inside the window
                    int r2 = (int)my_rel;
#pragma unroll
                    for (int k = 0; k < kIpt; k++) {
                        if (cnt[k] > 0) {
                            const unsigned d = (unsigned)(r2 - w0);
                            if (d < (unsigned)kWindow) atomicOr(&H[d >> 5].x, 1u << (d & 31));
                            r2 += cnt[k];
                        }
                    }
                }
                __syncthreads();
                {   // rank_base[word] = carried + heads in earlier words (2 words per thread)
                    const int c0 = __popc(H[2 * tid].x), c1 = __popc(H[2 * tid + 1].x);
                    int wsum;
                    const int ex = warp_excl_scan(c0 + c1, wsum);
                    if (lane == 0) sm.warp_heads[warp] = wsum;
                    __syncthreads();
                    int before = carried, all = 0;
#pragma unroll
                    for (int w = 0; w < kWarps; w++) {
                        const int c = sm.warp_heads[w];
                        before += (w < warp) ? c : 0;
                        all += c;
                    }
                    H[2 * tid].y = (unsigned)(before + ex);
                    H[2 * tid + 1].y = (unsigned)(before + ex + c0);
                    carried += all;
                }
                __syncthreads();
                const int wend = min(total - w0, kWindow);
                const long long obase = (long long)tile_base + w0;
                for (int q = tid; q < wend; q += kThreads) {
                    const uint2 hw = H[q >> 5];
                    const int r = (int)hw.y + __popc(hw.x & lane_le) - 1;
                    int rrel;
                    long long src;
                    sm.rec.get(r, rrel, src);
                    src += w0 + q - rrel;
                    const long long o = obase + q;
                    if (o < p.capacity) {
                        double t;
                        int s;
                        if (SMALL) {
                            t = s_start[(int)src];
                            s = s_sn[(int)src];
                        } else {
                            t = staged ? s_start[(int)(src - span_lo)] : __ldg(p.start + src);
                            s = __ldg(p.sn + src);
                        }
                        __stcs(p.time_out + o, t);   // time = fragment.start  (media-map.js:48)
                        __stcs(p.sn_out + o, s);     // sn   = fragment.sn     (media-map.js:46)
                    }
                }
                __syncthreads();
                H[2 * tid].x = 0u;        // this buffer is next used two windows from now
                H[2 * tid + 1].x = 0u;
                hbuf ^= 1u;
            }
        }

        if (Scan::tile_done(p.ctr, p.n_tiles)) Scan::finish_batch(p.ctr, p.result, p.epoch);
    }
    // never leave with a bulk copy still landing in this CTA's shared memory
    if (stage_pending) mbar_wait(&sm.mbar, mbar_phase);
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
constexpr int kFWarps = kFThreads / 32;
using FScan = TileScan<kFThreads>;
static_assert(kFThreads == kFilterTileQueries, "filter tile = one query per thread");
}  // namespace

__global__ void __launch_bounds__(kFThreads)
resolve_filter_kernel(const ResolveParams p)
{
    __shared__ FScan::Smem scan;
    __shared__ unsigned long long warp_cnt[kFWarps];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lane_lt = (1u << lane) - 1u;

    for (;;) {
        const int tile = FScan::next_tile(scan, p.ctr, p.epoch);
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
            for (int i0 = 0; i0 < jlen; i0 += 32) {
                const int i = i0 + lane;
                bool keep = false;
                if (i < jlen) {
                    const double s = __ldg(p.start + joff + i);
                    keep = (jbg <= s) && (s <= jub);
                }
                c += __popc(__ballot_sync(0xffffffffu, keep));
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
            FScan::exclusive_prefix(scan, p.tile_status, p.epoch, tile, tile_total);
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
                for (int i0 = 0; i0 < jlen; i0 += 32) {
                    const int i = i0 + lane;
                    bool keep = false;
                    double s = 0.0;
                    if (i < jlen) {
                        s = __ldg(p.start + joff + i);
                        keep = (jbg <= s) && (s <= jub);
                    }
                    const unsigned ball = __ballot_sync(0xffffffffu, keep);
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

        if (FScan::tile_done(p.ctr, p.n_tiles)) FScan::finish_batch(p.ctr, p.result, p.epoch);
    }
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

// =========================================================================================
//  upload-time monotonicity proof: one warp per track
// =========================================================================================
__global__ void __launch_bounds__(256)
monotone_check_kernel(const double *__restrict__ start, TrackDesc *desc, int n_tracks)
{
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    for (int t = blockIdx.x * wpb + (threadIdx.x >> 5); t < n_tracks; t += gridDim.x * wpb) {
        TrackDesc d = desc[t];
        if ((d.flags & kTrackStateMask) != HLSMAP_TRACK_PARSED) continue;
        bool bad = false;
        for (int i = lane; i < d.len; i += 32) {
            const double s = start[d.off + i];
            bad |= (s != s);
            if (i > 0) bad |= (s < start[d.off + i - 1]);
        }
        const unsigned any_bad = __ballot_sync(0xffffffffu, bad);
        if (lane == 0 && !any_bad) desc[t].flags = d.flags | kTrackMonotone;
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

int resolve_search_max_grid(size_t smem, bool small)
{
    int per_sm = 0;
    cudaError_t e;
    if (small) {
        cudaFuncSetAttribute(resolve_search_kernel<true>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resolve_search_kernel<true>,
                                                          kThreads, smem);
    } else {
        cudaFuncSetAttribute(resolve_search_kernel<false>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resolve_search_kernel<false>,
                                                          kThreads, smem);
    }
    if (e != cudaSuccess || per_sm < 1) per_sm = 1;
    return per_sm * sm_count();
}

int resolve_filter_max_grid()
{
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resolve_filter_kernel, kFThreads,
                                                      0) != cudaSuccess || per_sm < 1)
        per_sm = 1;
    return per_sm * sm_count();
}

cudaError_t launch_resolve_search(const ResolveParams &p, int grid, size_t smem, cudaStream_t s)
{
    if (p.small) resolve_search_kernel<true><<<grid, kThreads, smem, s>>>(p);
    else resolve_search_kernel<false><<<grid, kThreads, smem, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_resolve_filter(const ResolveParams &p, int grid, cudaStream_t s)
{
    resolve_filter_kernel<<<grid, kFThreads, 0, s>>>(p);
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

cudaError_t launch_monotone_check(const double *start, TrackDesc *desc, int n_tracks,
                                  cudaStream_t s)
{
    if (n_tracks <= 0) return cudaSuccess;
    long long blocks = ((long long)n_tracks + 7) / 8;
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    monotone_check_kernel<<<(int)blocks, 256, 0, s>>>(start, desc, n_tracks);
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

}  // namespace hlsmap

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

#include "scan.cuh"

namespace hlsmap {

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kIpt = 4;                         // queries per thread
constexpr int kTile = kThreads * kIpt;          // == kSearchTileQueries
constexpr int kWarpQueries = 32 * kIpt;         // 128
constexpr int kMaxWarpSpan = 2048;              // outputs a warp expands via the head bitmap
static_assert(kTile == kSearchTileQueries, "tile size mismatch");

using Scan = TileScan<kThreads>;

struct __align__(16) ExpRec {  // one query with >= 1 hit, in query order
    long long src;             // first kept fragment (absolute index into start[]/sn[])
    int rel;                   // first output slot relative to the warp's output run
    int cnt;
};

struct WarpExp {
    ExpRec rec[kWarpQueries];
    unsigned heads[kMaxWarpSpan / 32];
};

struct SearchSmem {
    Scan::Smem scan;
    unsigned long long mbar;
    long long red_min[kWarps];
    long long red_max[kWarps];
    unsigned long long warp_cnt[kWarps];
    WarpExp wexp[kWarps];
};

__device__ __forceinline__ unsigned smem_u32(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}

// ---- TMA (bulk async copy) staging of a span of the start column -----------------------
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

// First i in [0, len) with  begin <= col(i)  (the filter's left test); len if none.
template <class Col>
__device__ __forceinline__ int lower_bound_begin(const Col &col, int len, double begin)
{
    int lo = 0, n = len;
    while (n > 0) {
        const int half = n >> 1;
        const bool keep_right = !(begin <= col(lo + half));
        lo = keep_right ? lo + half + 1 : lo;
        n = keep_right ? n - half - 1 : half;
    }
    return lo;
}

// First i in [lo, len) with  !(col(i) <= ub)  (the filter's right test fails); gallops
// from lo because a look-ahead window keeps only a handful of fragments.
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
    while (a < b) {
        const int m = (a + b) >> 1;
        if (col(m) <= ub) a = m + 1;
        else b = m;
    }
    return a;
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

}  // namespace

// =========================================================================================
//  search kernel
// =========================================================================================
__global__ void __launch_bounds__(kThreads)
resolve_search_kernel(const ResolveParams p)
{
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    double *s_start = reinterpret_cast<double *>(dyn_smem);
    __shared__ SearchSmem sm;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lane_le = 0xffffffffu >> (31 - lane);

    if (tid == 0) mbar_init(&sm.mbar, 1);
    for (int i = lane; i < kMaxWarpSpan / 32; i += 32) sm.wexp[warp].heads[i] = 0u;
    unsigned mbar_phase = 0;
    long long span_lo = 0, span_hi = 0;  // staged element range of start[]
    __syncthreads();

    if (p.stage_all) {
        span_lo = 0;
        span_hi = p.n_frag_padded;
        if (tid == 0) {
            const unsigned bytes = (unsigned)(span_hi * 8);
            mbar_expect_tx(&sm.mbar, bytes);
            for (unsigned o = 0; o < bytes; o += 32768u)
                bulk_g2s(dyn_smem + o, reinterpret_cast<const unsigned char *>(p.start) + o,
                         min(32768u, bytes - o), &sm.mbar);
        }
    }
    bool stage_pending = p.stage_all != 0;

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
        int len[kIpt], st[kIpt];
        int n_unp = 0, n_mis = 0;
        long long mn = 0x7fffffffffffffffll, mx = -1;
#pragma unroll
        for (int k = 0; k < kIpt; k++) {
            unsigned fl;
            st[k] = lookup_track(p.desc, p.playlist_base, p.n_playlists, pl[k], lv[k], off[k],
                                 len[k], fl);
            const bool in = q0 + k < p.nq;
            n_unp += (in && st[k] == HLSMAP_ST_UNPARSED);
            n_mis += (in && st[k] == HLSMAP_ST_MISSING);
            if (st[k] == HLSMAP_ST_OK && len[k] > 0) {
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
        bool staged;
        if (p.stage_all) {
            staged = true;
        } else {
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
                staged = true;
                span_lo = need_lo;
                span_hi = need_hi;
                if (tid == 0) {
                    const unsigned bytes = (unsigned)((span_hi - span_lo) * 8);
                    const unsigned char *src =
                        reinterpret_cast<const unsigned char *>(p.start + span_lo);
                    mbar_expect_tx(&sm.mbar, bytes);
                    for (unsigned o = 0; o < bytes; o += 32768u)
                        bulk_g2s(dyn_smem + o, src + o, min(32768u, bytes - o), &sm.mbar);
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

        // ---- bounds: [lo, hi) is exactly what the filter keeps ---------------------------
        int lo[kIpt], cnt[kIpt];
#pragma unroll
        for (int k = 0; k < kIpt; k++) {
            // one fp64 add, round-to-nearest, never contracted: JS `beginTime + duration`
            const double ub = __dadd_rn(bg[k], du[k]);
            int l, h;
            if (staged) {
                const SmemCol col{s_start + (int)(off[k] - span_lo)};
                l = lower_bound_begin(col, len[k], bg[k]);
                h = upper_bound_from(col, l, len[k], ub);
            } else {
                const GmemCol col{p.start + off[k]};
                l = lower_bound_begin(col, len[k], bg[k]);
                h = upper_bound_from(col, l, len[k], ub);
            }
            lo[k] = l;
            cnt[k] = h - l;
        }

        // ---- scan counts: thread -> warp -> CTA -> batch ---------------------------------
        const int tsum = cnt[0] + cnt[1] + cnt[2] + cnt[3];
        int wtotal;
        const int texcl = warp_excl_scan(tsum, wtotal);
        if (lane == 0) sm.warp_cnt[warp] = (unsigned long long)wtotal;
        __syncthreads();
        unsigned long long wbase = 0, tile_total = 0;
#pragma unroll
        for (int w = 0; w < kWarps; w++) {
            const unsigned long long c = sm.warp_cnt[w];
            wbase += (w < warp) ? c : 0ull;
            tile_total += c;
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
        const unsigned long long tile_base =
            Scan::exclusive_prefix(sm.scan, p.tile_status, p.epoch, tile, tile_total);

        // ---- offsets[q] ------------------------------------------------------------------
        const unsigned long long my_base = tile_base + wbase + (unsigned long long)texcl;
        {
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
            const unsigned long long total = tile_base + tile_total;
            p.offsets[p.nq] = (int)total;
            p.result->total_hits = (long long)total;
            p.result->overflow = total > 0x7fffffffull ? 1 : 0;
        }

        // ---- expansion: the warp's hits are one contiguous output run --------------------
        if (p.sn_out != nullptr && wtotal > 0) {
            WarpExp &W = sm.wexp[warp];
            const int nz = (cnt[0] > 0) + (cnt[1] > 0) + (cnt[2] > 0) + (cnt[3] > 0);
            int nrec;
            int rank = warp_excl_scan(nz, nrec);
            const bool bitmap = wtotal <= kMaxWarpSpan;
            int rel = texcl;
#pragma unroll
            for (int k = 0; k < kIpt; k++) {
                if (cnt[k] > 0) {
                    ExpRec r;
                    r.src = off[k] + lo[k];
                    r.rel = rel;
                    r.cnt = cnt[k];
                    W.rec[rank] = r;
                    if (bitmap) atomicOr(&W.heads[rel >> 5], 1u << (rel & 31));
                    rank++;
                    rel += cnt[k];
                }
            }
            __syncwarp();
            const long long gbase = (long long)(tile_base + wbase);
            if (bitmap) {
                int carried = 0;
                for (int c0 = 0; c0 < wtotal; c0 += 32) {
                    const unsigned m = W.heads[c0 >> 5];
                    const int q = c0 + lane;
                    const int r = carried + __popc(m & lane_le) - 1;
                    carried += __popc(m);
                    if (q < wtotal) {
                        const ExpRec rc = W.rec[r];
                        const long long src = rc.src + (q - rc.rel);
                        const long long o = gbase + q;
                        if (o < p.capacity) {
                            const double t = staged ? s_start[(int)(src - span_lo)]
                                                    : __ldg(p.start + src);
                            __stcs(p.time_out + o, t);            // time = fragment.start (:48)
                            __stcs(p.sn_out + o, __ldg(p.sn + src));  // sn = fragment.sn (:46)
                        }
                    }
                    __syncwarp();
                    if (lane == 0) W.heads[c0 >> 5] = 0u;
                }
            } else {
                for (int r = 0; r < nrec; r++) {
                    const ExpRec rc = W.rec[r];
                    for (int j = lane; j < rc.cnt; j += 32) {
                        const long long o = gbase + rc.rel + j;
                        if (o < p.capacity) {
                            p.time_out[o] = __ldg(p.start + rc.src + j);
                            p.sn_out[o] = __ldg(p.sn + rc.src + j);
                        }
                    }
                }
            }
            __syncwarp();
        }

        if (Scan::tile_done(p.ctr, p.n_tiles)) Scan::finish_batch(p.ctr, p.result, p.epoch);
    }
    // never leave with a bulk copy still landing in this CTA's shared memory
    if (stage_pending) mbar_wait(&sm.mbar, mbar_phase);
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
                const int l = lower_bound_begin(col, len, bg);
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

int resolve_search_max_grid(size_t smem)
{
    int per_sm = 0;
    cudaFuncSetAttribute(resolve_search_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)smem);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resolve_search_kernel, kThreads,
                                                      smem) != cudaSuccess || per_sm < 1)
        per_sm = 1;
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
    resolve_search_kernel<<<grid, kThreads, smem, s>>>(p);
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

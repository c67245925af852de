// views.cu -- SegmentView / TrackView value-type operations over SoA columns (sm_100a).
//
// A view is (level, urlId, sn[, time]) (segment-view.js:22-26, track-view.js:3-6).  These
// are pure streaming kernels: every thread moves 4 views with 128-bit loads and stores,
// so each is bounded by HBM bandwidth (12 B in / 12 B out per view for pack).
#include "scan.cuh"

namespace hlsmap {

namespace {

constexpr int kVThreads = 256;

__device__ __forceinline__ int view_grid_stride_begin(long long &i4, long long &stride4)
{
    i4 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    stride4 = (long long)gridDim.x * blockDim.x;
    return 0;
}

inline int view_blocks(long long n, int sm)
{
    long long b = (n / 4 + kVThreads) / kVThreads;
    const long long cap = (long long)sm * 16;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace

// toArrayBuffer: Uint32Array([level, urlId, sn]) (segment-view.js:59-61).  ToUint32 of an
// int32 is its two's-complement bit pattern, so this is a 3-way interleave.
__global__ void __launch_bounds__(kVThreads)
view_pack_kernel(long long n, const int *__restrict__ level, const int *__restrict__ url_id,
                 const int *__restrict__ sn, unsigned *__restrict__ wire, int vec)
{
    long long g, stride;
    view_grid_stride_begin(g, stride);
    const long long n4 = vec ? n / 4 : 0;
    for (long long i = g; i < n4; i += stride) {
        const int4 l = __ldcs(reinterpret_cast<const int4 *>(level) + i);
        const int4 u = __ldcs(reinterpret_cast<const int4 *>(url_id) + i);
        const int4 s = __ldcs(reinterpret_cast<const int4 *>(sn) + i);
        uint4 *w = reinterpret_cast<uint4 *>(wire) + 3 * i;
        __stcs(w + 0, make_uint4((unsigned)l.x, (unsigned)u.x, (unsigned)s.x, (unsigned)l.y));
        __stcs(w + 1, make_uint4((unsigned)u.y, (unsigned)s.y, (unsigned)l.z, (unsigned)u.z));
        __stcs(w + 2, make_uint4((unsigned)s.z, (unsigned)l.w, (unsigned)u.w, (unsigned)s.w));
    }
    for (long long i = n4 * 4 + g; i < n; i += stride) {
        wire[3 * i + 0] = (unsigned)level[i];
        wire[3 * i + 1] = (unsigned)url_id[i];
        wire[3 * i + 2] = (unsigned)sn[i];
    }
}

// fromArrayBuffer: first three u32 -> level, urlId, sn (segment-view.js:9-17).
__global__ void __launch_bounds__(kVThreads)
view_unpack_kernel(long long n, const unsigned *__restrict__ wire, unsigned *__restrict__ level,
                   unsigned *__restrict__ url_id, unsigned *__restrict__ sn, int vec)
{
    long long g, stride;
    view_grid_stride_begin(g, stride);
    const long long n4 = vec ? n / 4 : 0;
    for (long long i = g; i < n4; i += stride) {
        const uint4 *w = reinterpret_cast<const uint4 *>(wire) + 3 * i;
        const uint4 a = __ldcs(w + 0), b = __ldcs(w + 1), c = __ldcs(w + 2);
        __stcs(reinterpret_cast<uint4 *>(level) + i, make_uint4(a.x, a.w, b.z, c.y));
        __stcs(reinterpret_cast<uint4 *>(url_id) + i, make_uint4(a.y, b.x, b.w, c.z));
        __stcs(reinterpret_cast<uint4 *>(sn) + i, make_uint4(a.z, b.y, c.x, c.w));
    }
    for (long long i = n4 * 4 + g; i < n; i += stride) {
        level[i] = wire[3 * i + 0];
        url_id[i] = wire[3 * i + 1];
        sn[i] = wire[3 * i + 2];
    }
}

// isEqual: sn === sn' && level === level' && urlId === urlId' (segment-view.js:33-39,
// track-view.js:19-24); time is not compared.
__global__ void __launch_bounds__(kVThreads)
view_is_equal_kernel(long long n, const int *__restrict__ al, const int *__restrict__ au,
                     const int *__restrict__ as, const int *__restrict__ bl,
                     const int *__restrict__ bu, const int *__restrict__ bs,
                     unsigned char *__restrict__ out, int vec)
{
    long long g, stride;
    view_grid_stride_begin(g, stride);
    const long long n4 = vec ? n / 4 : 0;
    for (long long i = g; i < n4; i += stride) {
        const int4 a0 = __ldcs(reinterpret_cast<const int4 *>(al) + i);
        const int4 a1 = __ldcs(reinterpret_cast<const int4 *>(au) + i);
        const int4 a2 = __ldcs(reinterpret_cast<const int4 *>(as) + i);
        const int4 b0 = __ldcs(reinterpret_cast<const int4 *>(bl) + i);
        const int4 b1 = __ldcs(reinterpret_cast<const int4 *>(bu) + i);
        const int4 b2 = __ldcs(reinterpret_cast<const int4 *>(bs) + i);
        uchar4 r;
        r.x = (a2.x == b2.x) && (a0.x == b0.x) && (a1.x == b1.x);
        r.y = (a2.y == b2.y) && (a0.y == b0.y) && (a1.y == b1.y);
        r.z = (a2.z == b2.z) && (a0.z == b0.z) && (a1.z == b1.z);
        r.w = (a2.w == b2.w) && (a0.w == b0.w) && (a1.w == b1.w);
        __stcs(reinterpret_cast<uchar4 *>(out) + i, r);
    }
    for (long long i = n4 * 4 + g; i < n; i += stride)
        out[i] = (as[i] == bs[i]) && (al[i] == bl[i]) && (au[i] == bu[i]);
}

// isInTrack (segment-view.js:45-47): TrackView equality of the view's track.
__global__ void __launch_bounds__(kVThreads)
view_is_in_track_kernel(long long n, const int *__restrict__ al, const int *__restrict__ au,
                        const int *__restrict__ tl, const int *__restrict__ tu,
                        unsigned char *__restrict__ out, int vec)
{
    long long g, stride;
    view_grid_stride_begin(g, stride);
    const long long n4 = vec ? n / 4 : 0;
    for (long long i = g; i < n4; i += stride) {
        const int4 a0 = __ldcs(reinterpret_cast<const int4 *>(al) + i);
        const int4 a1 = __ldcs(reinterpret_cast<const int4 *>(au) + i);
        const int4 b0 = __ldcs(reinterpret_cast<const int4 *>(tl) + i);
        const int4 b1 = __ldcs(reinterpret_cast<const int4 *>(tu) + i);
        uchar4 r;
        r.x = (a0.x == b0.x) && (a1.x == b1.x);
        r.y = (a0.y == b0.y) && (a1.y == b1.y);
        r.z = (a0.z == b0.z) && (a1.z == b1.z);
        r.w = (a0.w == b0.w) && (a1.w == b1.w);
        __stcs(reinterpret_cast<uchar4 *>(out) + i, r);
    }
    for (long long i = n4 * 4 + g; i < n; i += stride)
        out[i] = (al[i] == tl[i]) && (au[i] == tu[i]);
}

// toArrayBuffer -> fromArrayBuffer -> isEqual(original), parse fused from registers.
// The parsed side holds Numbers in [0, 2^32) while the original holds int32 Numbers, so
// a field survives the trip iff its uint32 reading equals its int32 value.
__device__ __forceinline__ unsigned char rt_equal(int l, int u, int s)
{
    return ((long long)(unsigned)s == (long long)s) && ((long long)(unsigned)l == (long long)l) &&
           ((long long)(unsigned)u == (long long)u);
}

__global__ void __launch_bounds__(kVThreads)
view_roundtrip_kernel(long long n, const int *__restrict__ level, const int *__restrict__ url_id,
                      const int *__restrict__ sn, unsigned *__restrict__ wire,
                      unsigned char *__restrict__ equal, int vec)
{
    long long g, stride;
    view_grid_stride_begin(g, stride);
    const long long n4 = vec ? n / 4 : 0;
    for (long long i = g; i < n4; i += stride) {
        const int4 l = __ldcs(reinterpret_cast<const int4 *>(level) + i);
        const int4 u = __ldcs(reinterpret_cast<const int4 *>(url_id) + i);
        const int4 s = __ldcs(reinterpret_cast<const int4 *>(sn) + i);
        uint4 *w = reinterpret_cast<uint4 *>(wire) + 3 * i;
        __stcs(w + 0, make_uint4((unsigned)l.x, (unsigned)u.x, (unsigned)s.x, (unsigned)l.y));
        __stcs(w + 1, make_uint4((unsigned)u.y, (unsigned)s.y, (unsigned)l.z, (unsigned)u.z));
        __stcs(w + 2, make_uint4((unsigned)s.z, (unsigned)l.w, (unsigned)u.w, (unsigned)s.w));
        uchar4 r;
        r.x = rt_equal(l.x, u.x, s.x);
        r.y = rt_equal(l.y, u.y, s.y);
        r.z = rt_equal(l.z, u.z, s.z);
        r.w = rt_equal(l.w, u.w, s.w);
        __stcs(reinterpret_cast<uchar4 *>(equal) + i, r);
    }
    for (long long i = n4 * 4 + g; i < n; i += stride) {
        wire[3 * i + 0] = (unsigned)level[i];
        wire[3 * i + 1] = (unsigned)url_id[i];
        wire[3 * i + 2] = (unsigned)sn[i];
        equal[i] = rt_equal(level[i], url_id[i], sn[i]);
    }
}

// ---- viewToString: "L{level}U{urlId}S{sn}" (segment-view.js:52-54, track-view.js:11-13) --
namespace {

constexpr int kFmtThreads = 256;
constexpr int kFmtIpt = 4;
constexpr int kFmtMaxChars = 36;  // 3 letters + 3 x "-2147483648"
using FmtScan = TileScan<kFmtThreads>;
static_assert(kFmtThreads * kFmtIpt == kFormatTileViews, "format tile size mismatch");

__device__ __forceinline__ int dec_len(int v)
{
    unsigned m = v < 0 ? 0u - (unsigned)v : (unsigned)v;
    int d = 1;
    d += m >= 10u;
    d += m >= 100u;
    d += m >= 1000u;
    d += m >= 10000u;
    d += m >= 100000u;
    d += m >= 1000000u;
    d += m >= 10000000u;
    d += m >= 100000000u;
    d += m >= 1000000000u;
    return d + (v < 0);
}

// JS Number -> String of an int32: plain decimal, '-' prefix when negative.
__device__ __forceinline__ int put_dec(char *dst, int v)
{
    const int n = dec_len(v);
    unsigned m = v < 0 ? 0u - (unsigned)v : (unsigned)v;
    for (int i = n - 1; i >= (v < 0); i--) {
        dst[i] = (char)('0' + m % 10u);
        m /= 10u;
    }
    if (v < 0) dst[0] = '-';
    return n;
}

}  // namespace

__global__ void __launch_bounds__(kFmtThreads)
view_format_kernel(const FormatParams p)
{
    __shared__ FmtScan::Smem scan;
    __shared__ unsigned long long warp_cnt[kFmtThreads / 32];
    __shared__ __align__(16) char buf[kFormatTileViews * kFmtMaxChars + 16];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    FmtScan::acquire_set(p.ctr, p.tslot);
    for (;;) {
        const int tile = FmtScan::next_tile(scan, p.ctr, p.tslot);
        if (tile >= p.n_tiles) break;
        const long long v0 = (long long)tile * kFormatTileViews + (long long)tid * kFmtIpt;
        int l[kFmtIpt], u[kFmtIpt], s[kFmtIpt], len[kFmtIpt];
        int tsum = 0;
#pragma unroll
        for (int k = 0; k < kFmtIpt; k++) {
            const bool in = v0 + k < p.n;
            l[k] = in ? p.level[v0 + k] : 0;
            u[k] = in ? p.url_id[v0 + k] : 0;
            s[k] = (in && p.with_sn) ? p.sn[v0 + k] : 0;
            len[k] = in ? 2 + dec_len(l[k]) + dec_len(u[k]) + (p.with_sn ? 1 + dec_len(s[k]) : 0) : 0;
            tsum += len[k];
        }
        int x = tsum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        const int wtotal = __shfl_sync(0xffffffffu, x, 31);
        const int texcl = x - tsum;
        if (lane == 0) warp_cnt[warp] = (unsigned long long)wtotal;
        __syncthreads();
        unsigned long long wbase = 0, tile_total = 0;
#pragma unroll
        for (int w = 0; w < kFmtThreads / 32; w++) {
            const unsigned long long c = warp_cnt[w];
            wbase += (w < warp) ? c : 0ull;
            tile_total += c;
        }
        const unsigned long long tile_base =
            FmtScan::exclusive_prefix(scan, p.tile_status, p.epoch, tile, tile_total);

        // the tile's characters sit in shared memory at the same offset modulo 16 as their place in
        // global memory, so the copy out below moves 16 bytes per thread with aligned accesses
        const int pad = (int)((reinterpret_cast<unsigned long long>(p.chars) + tile_base) & 15ull);
        int at = (int)wbase + texcl;  // tile-relative byte offset of this thread's first view
#pragma unroll
        for (int k = 0; k < kFmtIpt; k++) {
            if (v0 + k < p.n) {
                p.offsets[v0 + k] = (int)(tile_base + (unsigned long long)at);
                char *d = buf + pad + at;
                *d++ = 'L';
                d += put_dec(d, l[k]);
                *d++ = 'U';
                d += put_dec(d, u[k]);
                if (p.with_sn) {
                    *d++ = 'S';
                    d += put_dec(d, s[k]);
                }
                at += len[k];
            }
        }
        if (tile == p.n_tiles - 1 && tid == 0) {
            const unsigned long long total = tile_base + tile_total;
            p.offsets[p.n] = (int)total;
            p.result->total_hits = (long long)total;
            p.result->overflow = total > 0x7fffffffull ? 1 : 0;
        }
        __syncthreads();
        {   // coalesced copy of the tile's characters: bytes up to the first 16-byte boundary, whole
            // 16-byte pieces, bytes after the last boundary
            const int total = (int)tile_total;
            char *g = p.chars + (long long)tile_base - pad;          // 16-byte aligned
            if ((long long)tile_base + total <= p.capacity) {
                const int first = pad ? 16 : 0, last = (pad + total) & ~15;
                if (last > first) {
                    for (int i = first + 16 * tid; i < last; i += 16 * kFmtThreads)
                        *reinterpret_cast<uint4 *>(g + i) = *reinterpret_cast<const uint4 *>(buf + i);
                    if (tid < 16 && pad && pad + tid < 16) g[pad + tid] = buf[pad + tid];
                    if (tid >= 32 && tid < 48 && last + (tid - 32) < pad + total) g[last + tid - 32] = buf[last + tid - 32];
                } else {
                    for (int i = pad + tid; i < pad + total; i += kFmtThreads) g[i] = buf[i];
                }
            } else {
                for (int i = tid; i < total; i += kFmtThreads)
                    if ((long long)tile_base + i < p.capacity) g[pad + i] = buf[pad + i];
            }
        }
        if (FmtScan::tile_done(p.ctr, p.n_tiles)) FmtScan::finish_batch(p.ctr, p.result, p.tslot);
    }
    FmtScan::leave_set(p.ctr, p.tslot);
}

// ---- the same in three launches, none of which waits for another CTA -----------------------
// (1) characters per 1024-view tile, (2) one CTA scans the tile sums, (3) render with known bases.
// The columns are read twice (12 B a view more); in exchange no tile waits in a look-back for the
// slowest tile ahead of it.
__global__ void __launch_bounds__(kFmtThreads)
view_format_measure_kernel(const FormatParams p)
{
    __shared__ unsigned warp_sum[kFmtThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long v0 = (long long)blockIdx.x * kFormatTileViews + (long long)tid * kFmtIpt;
    int x = 0;
#pragma unroll
    for (int k = 0; k < kFmtIpt; k++)
        if (v0 + k < p.n)
            x += 2 + dec_len(p.level[v0 + k]) + dec_len(p.url_id[v0 + k]) + (p.with_sn ? 1 + dec_len(p.sn[v0 + k]) : 0);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) x += __shfl_xor_sync(0xffffffffu, x, d);
    if (lane == 0) warp_sum[warp] = (unsigned)x;
    __syncthreads();
    if (tid == 0) {
        unsigned t = 0;
#pragma unroll
        for (int w = 0; w < kFmtThreads / 32; w++) t += warp_sum[w];
        p.block_sum[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(1024)
tile_sum_scan_kernel(const unsigned long long *sum, unsigned long long *base, int n_tiles, int *offsets_end,
                     DeviceResult *result)
{
    __shared__ unsigned long long wsum[32];
    __shared__ unsigned long long carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int b0 = 0; b0 < n_tiles; b0 += 1024) {
        const int b = b0 + tid;
        const unsigned long long v = b < n_tiles ? sum[b] : 0ull;
        unsigned long long x = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) wsum[warp] = x;
        __syncthreads();
        unsigned long long before = carry, all = 0;
        for (int w = 0; w < 32; w++) {
            const unsigned long long c = wsum[w];
            before += w < warp ? c : 0ull;
            all += c;
        }
        if (b < n_tiles) base[b] = before + x - v;
        __syncthreads();
        if (tid == 0) carry += all;
        __syncthreads();
    }
    if (tid == 0) {
        const unsigned long long total = carry;
        *offsets_end = (int)total;
        result->total_hits = (long long)total;
        result->n_unparsed = 0;
        result->n_missing = 0;
        result->overflow = total > 0x7fffffffull ? 1 : 0;
    }
}

__global__ void __launch_bounds__(kFmtThreads)
view_format_render_kernel(const FormatParams p)
{
    __shared__ unsigned warp_cnt[kFmtThreads / 32];
    __shared__ __align__(16) char buf[kFormatTileViews * kFmtMaxChars + 16];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long v0 = (long long)blockIdx.x * kFormatTileViews + (long long)tid * kFmtIpt;
    int l[kFmtIpt], u[kFmtIpt], s[kFmtIpt], len[kFmtIpt];
    int tsum = 0;
#pragma unroll
    for (int k = 0; k < kFmtIpt; k++) {
        const bool in = v0 + k < p.n;
        l[k] = in ? p.level[v0 + k] : 0;
        u[k] = in ? p.url_id[v0 + k] : 0;
        s[k] = (in && p.with_sn) ? p.sn[v0 + k] : 0;
        len[k] = in ? 2 + dec_len(l[k]) + dec_len(u[k]) + (p.with_sn ? 1 + dec_len(s[k]) : 0) : 0;
        tsum += len[k];
    }
    int x = tsum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
    }
    if (lane == 31) warp_cnt[warp] = (unsigned)x;
    __syncthreads();
    unsigned wbase = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < kFmtThreads / 32; w++) {
        const unsigned c = warp_cnt[w];
        wbase += (w < warp) ? c : 0u;
        tile_total += c;
    }
    const unsigned long long tile_base = p.block_base[blockIdx.x];
    const int pad = (int)((reinterpret_cast<unsigned long long>(p.chars) + tile_base) & 15ull);
    int at = (int)wbase + x - tsum;
#pragma unroll
    for (int k = 0; k < kFmtIpt; k++) {
        if (v0 + k < p.n) {
            p.offsets[v0 + k] = (int)(tile_base + (unsigned long long)at);
            char *d = buf + pad + at;
            *d++ = 'L';
            d += put_dec(d, l[k]);
            *d++ = 'U';
            d += put_dec(d, u[k]);
            if (p.with_sn) {
                *d++ = 'S';
                d += put_dec(d, s[k]);
            }
            at += len[k];
        }
    }
    __syncthreads();
    const int total = (int)tile_total;
    char *g = p.chars + (long long)tile_base - pad;          // 16-byte aligned
    if ((long long)tile_base + total <= p.capacity) {
        const int first = pad ? 16 : 0, last = (pad + total) & ~15;
        if (last > first) {
            for (int i = first + 16 * tid; i < last; i += 16 * kFmtThreads)
                *reinterpret_cast<uint4 *>(g + i) = *reinterpret_cast<const uint4 *>(buf + i);
            if (tid < 16 && pad && pad + tid < 16) g[pad + tid] = buf[pad + tid];
            if (tid >= 32 && tid < 48 && last + (tid - 32) < pad + total) g[last + tid - 32] = buf[last + tid - 32];
        } else {
            for (int i = pad + tid; i < pad + total; i += kFmtThreads) g[i] = buf[i];
        }
    } else {
        for (int i = tid; i < total; i += kFmtThreads)
            if ((long long)tile_base + i < p.capacity) g[pad + i] = buf[pad + i];
    }
}

cudaError_t launch_view_format_split(const FormatParams &p, cudaStream_t s)
{
    if (p.n_tiles < 1) return cudaSuccess;
    view_format_measure_kernel<<<p.n_tiles, kFmtThreads, 0, s>>>(p);
    tile_sum_scan_kernel<<<1, 1024, 0, s>>>(p.block_sum, p.block_base, p.n_tiles, p.offsets + p.n, p.result);
    view_format_render_kernel<<<p.n_tiles, kFmtThreads, 0, s>>>(p);
    return cudaGetLastError();
}

// ---- launchers ---------------------------------------------------------------------------
static int views_sm_count()
{
    static int n = 0;
    if (!n) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

cudaError_t launch_view_pack(long long n, const int *level, const int *url_id, const int *sn,
                             unsigned *wire, cudaStream_t s)
{
    const int vec = aligned16(level) && aligned16(url_id) && aligned16(sn) && aligned16(wire);
    view_pack_kernel<<<view_blocks(n, views_sm_count()), kVThreads, 0, s>>>(n, level, url_id, sn,
                                                                           wire, vec);
    return cudaGetLastError();
}

cudaError_t launch_view_unpack(long long n, const unsigned *wire, unsigned *level,
                               unsigned *url_id, unsigned *sn, cudaStream_t s)
{
    const int vec = aligned16(level) && aligned16(url_id) && aligned16(sn) && aligned16(wire);
    view_unpack_kernel<<<view_blocks(n, views_sm_count()), kVThreads, 0, s>>>(n, wire, level,
                                                                             url_id, sn, vec);
    return cudaGetLastError();
}

cudaError_t launch_view_is_equal(long long n, const int *al, const int *au, const int *as,
                                 const int *bl, const int *bu, const int *bs, unsigned char *out,
                                 cudaStream_t s)
{
    const int vec = aligned16(al) && aligned16(au) && aligned16(as) && aligned16(bl) &&
                    aligned16(bu) && aligned16(bs) && ((reinterpret_cast<uintptr_t>(out) & 3u) == 0);
    view_is_equal_kernel<<<view_blocks(n, views_sm_count()), kVThreads, 0, s>>>(n, al, au, as, bl,
                                                                               bu, bs, out, vec);
    return cudaGetLastError();
}

cudaError_t launch_view_is_in_track(long long n, const int *al, const int *au, const int *tl,
                                    const int *tu, unsigned char *out, cudaStream_t s)
{
    const int vec = aligned16(al) && aligned16(au) && aligned16(tl) && aligned16(tu) &&
                    ((reinterpret_cast<uintptr_t>(out) & 3u) == 0);
    view_is_in_track_kernel<<<view_blocks(n, views_sm_count()), kVThreads, 0, s>>>(n, al, au, tl,
                                                                                  tu, out, vec);
    return cudaGetLastError();
}

cudaError_t launch_view_roundtrip(long long n, const int *level, const int *url_id, const int *sn,
                                  unsigned *wire, unsigned char *equal, cudaStream_t s)
{
    const int vec = aligned16(level) && aligned16(url_id) && aligned16(sn) && aligned16(wire) &&
                    ((reinterpret_cast<uintptr_t>(equal) & 3u) == 0);
    view_roundtrip_kernel<<<view_blocks(n, views_sm_count()), kVThreads, 0, s>>>(
        n, level, url_id, sn, wire, equal, vec);
    return cudaGetLastError();
}

int view_format_max_grid()
{
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, view_format_kernel, kFmtThreads,
                                                      0) != cudaSuccess || per_sm < 1)
        per_sm = 1;
    return per_sm * views_sm_count();
}

cudaError_t launch_view_format(const FormatParams &p, int grid, cudaStream_t s)
{
    view_format_kernel<<<grid, kFmtThreads, 0, s>>>(p);
    return cudaGetLastError();
}

}  // namespace hlsmap

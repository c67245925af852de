// json.cu -- JSON.stringify(segmentView) and new SegmentView(JSON.parse(text)), batched.
//
// SURVEY 8 row f3.  The reference serialises a view for transfer with JSON (test/segment-view.js:
// 5-11); what JSON.stringify walks is the object the constructor built (segment-view.js:22-26:
// `sn`, `trackView` -- a TrackView with `level`, `urlId`, track-view.js:6-9 -- then `time`), so the
// text is   {"sn":25,"trackView":{"level":1,"urlId":0},"time":12.5}
// with properties whose value is undefined left out, numbers printed by Number::toString and a
// non-finite `time` printed as null.  Columns in, one string per view out -- and back.
#include "common.cuh"
#include "decimal.cuh"
#include "dtoa.cuh"
#include "scan.cuh"

namespace hlsmap {

__device__ const unsigned long long g_pow10_table[kPow10TableLen][2] = {
#include "pow10_table.inc"
};

namespace {

constexpr int kThreads = 256;
constexpr int kMaxChars = 104;   // {"sn":-2147483648,"trackView":{"level":-2147483648,"urlId":-2147483648},"time":-1.2345678901234567e-308}
using Scan = TileScan<kThreads>;
static_assert(kThreads == kJsonTileViews, "one view per thread");

__device__ __forceinline__ int int_len(int v)
{
    unsigned m = v < 0 ? 0u - (unsigned)v : (unsigned)v;
    int d = 1;
    while (m >= 10u) {
        m /= 10u;
        d++;
    }
    return d + (v < 0);
}

__device__ __forceinline__ int put_int(char *dst, int v)
{
    const int n = int_len(v);
    unsigned m = v < 0 ? 0u - (unsigned)v : (unsigned)v;
    for (int i = n - 1; i >= (v < 0); i--) {
        dst[i] = (char)('0' + m % 10u);
        m /= 10u;
    }
    if (v < 0) dst[0] = '-';
    return n;
}

__device__ __forceinline__ int put_lit(char *dst, const char *lit, int n)
{
    for (int i = 0; i < n; i++) dst[i] = lit[i];
    return n;
}

}  // namespace

__global__ void __launch_bounds__(kThreads)
view_to_json_kernel(const JsonWriteParams p)
{
    __shared__ Scan::Smem scan;
    __shared__ unsigned long long warp_cnt[kThreads / 32];
    __shared__ __align__(16) char buf[kThreads * kMaxChars + 16];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    Scan::acquire_set(p.ctr, p.epoch);
    for (;;) {
        const int tile = Scan::next_tile(scan, p.ctr, p.tslot);
        if (tile >= p.n_tiles) break;
        const long long v = (long long)tile * kThreads + tid;
        const bool in = v < p.n;
        const int l = in ? p.level[v] : 0, u = in ? p.url_id[v] : 0, s = in ? p.sn[v] : 0;
        // time: 0 = property absent, 1 = null (NaN, +-Infinity), 2 = a number
        int tkind = 0, tlen = 0;
        bool tneg = false;
        Shortest sd;
        sd.digits = 0;
        sd.exp10 = 0;
        sd.ndigits = 1;
        if (in && p.time != nullptr) {
            const unsigned long long bits = (unsigned long long)__double_as_longlong(p.time[v]);
            const unsigned long long mag = bits & 0x7fffffffffffffffull;
            if (mag >= 0x7ff0000000000000ull) {
                tkind = 1;
                tlen = 4;
            } else {
                tkind = 2;
                if (mag == 0) {
                    tlen = 1;                        // both zeros print as "0"
                } else {
                    tneg = (bits >> 63) != 0;
                    sd = shortest_digits(mag, g_pow10_table);
                    tlen = js_number_length(sd) + (tneg ? 1 : 0);
                }
            }
        }
        // {"sn":S,"trackView":{"level":L,"urlId":U}[,"time":T]}
        const int len = in ? 6 + int_len(s) + 22 + int_len(l) + 9 + int_len(u) + 1 + (tkind ? 8 + tlen : 0) + 1 : 0;

        int x = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) warp_cnt[warp] = (unsigned long long)x;
        __syncthreads();
        unsigned long long wbase = 0, tile_total = 0;
#pragma unroll
        for (int w = 0; w < kThreads / 32; w++) {
            const unsigned long long c = warp_cnt[w];
            wbase += (w < warp) ? c : 0ull;
            tile_total += c;
        }
        const unsigned long long tile_base = Scan::exclusive_prefix(scan, p.tile_status, p.epoch, tile, tile_total);
        const int at = (int)wbase + x - len;
        // the tile's characters sit in shared memory at the same offset modulo 16 as their place in
        // global memory, so the copy out below can move 16 bytes per thread with aligned accesses
        const int pad = (int)((reinterpret_cast<unsigned long long>(p.chars) + tile_base) & 15ull);
        if (in) {
            p.offsets[v] = (int)(tile_base + (unsigned long long)at);
            char *d = buf + pad + at;
            d += put_lit(d, "{\"sn\":", 6);
            d += put_int(d, s);
            d += put_lit(d, ",\"trackView\":{\"level\":", 22);
            d += put_int(d, l);
            d += put_lit(d, ",\"urlId\":", 9);
            d += put_int(d, u);
            *d++ = '}';
            if (tkind) {
                d += put_lit(d, ",\"time\":", 8);
                if (tkind == 1) {
                    d += put_lit(d, "null", 4);
                } else if (sd.digits == 0) {
                    *d++ = '0';
                } else {
                    if (tneg) *d++ = '-';
                    d += js_number_write(sd, d);
                }
            }
            *d++ = '}';
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
                    for (int i = first + 16 * tid; i < last; i += 16 * kThreads)
                        *reinterpret_cast<uint4 *>(g + i) = *reinterpret_cast<const uint4 *>(buf + i);
                    if (tid < 16 && pad && pad + tid < 16) g[pad + tid] = buf[pad + tid];
                    if (tid >= 32 && tid < 48 && last + (tid - 32) < pad + total) g[last + tid - 32] = buf[last + tid - 32];
                } else {
                    for (int i = pad + tid; i < pad + total; i += kThreads) g[i] = buf[i];
                }
            } else {
                for (int i = tid; i < total; i += kThreads)
                    if ((long long)tile_base + i < p.capacity) g[pad + i] = buf[pad + i];
            }
        }
        if (Scan::tile_done(p.ctr, p.n_tiles)) Scan::finish_batch(p.ctr, p.result, p.tslot);
    }
}

// ---- parse: the text JSON.stringify(segmentView) produces -> columns ------------------------
namespace {

// The characters of a CTA's 256 texts: their shared-memory copy when it was staged (positions are
// absolute offsets into the batch's character buffer), global memory otherwise.
struct Text {
    const char *g;
    const char *s;      // shared copy of [lo, hi), or nullptr
    long long lo;
    __device__ __forceinline__ int operator()(long long p) const
    {
        return (int)(unsigned char)(s ? s[p - lo] : g[p]);
    }
};

__device__ __forceinline__ bool eat(const Text &t, long long &q, long long lim, const char *lit, int n)
{
    if (q + n > lim) return false;
    for (int i = 0; i < n; i++)
        if (t(q + i) != (int)(unsigned char)lit[i]) return false;
    q += n;
    return true;
}

// A JSON number that must be an int32 (sn, level, urlId columns are int32).  What JSON.stringify
// prints for an int32 -- an optional '-' and up to ten digits -- is read directly; anything else
// (a fraction, an exponent, more digits) goes through the general number scanner.
__device__ __forceinline__ int eat_int(const Text &t, long long &q, long long lim, int *out)
{
    {
        long long r = q;
        const bool neg = r < lim && t(r) == '-';
        if (neg) r++;
        const long long d0 = r;
        long long val = 0;
        while (r < lim && r - d0 < 10) {
            const int c = t(r);
            if (c < '0' || c > '9') break;
            val = val * 10 + (c - '0');
            r++;
        }
        const int nd = (int)(r - d0);
        const int nx = r < lim ? t(r) : 0;
        const bool more = (nx >= '0' && nx <= '9') || nx == '.' || nx == 'e' || nx == 'E';
        if (nd > 0 && !more && !(nd > 1 && t(d0) == '0') && val <= (neg ? 2147483648ll : 2147483647ll)) {
            *out = (int)(neg ? -val : val);
            q = r;
            return HLSMAP_JSON_OK;
        }
    }
    const JsonNumber n = scan_json_number(t, q, lim);
    if (n.length == 0) return HLSMAP_JSON_MALFORMED;
    q += n.length;
    double x = 0.0;
    if (!json_number_to_double(n, &x)) return HLSMAP_JSON_UNSUPPORTED;
    if (!(x >= -2147483648.0 && x <= 2147483647.0) || x != (double)(int)x) return HLSMAP_JSON_NOT_INT32;
    *out = (int)x;
    return HLSMAP_JSON_OK;
}

__constant__ double c_json_pow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                        1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

// `time`: digits [. digits] with at most 15 significant digits and no exponent is one exact IEEE
// division (both operands are exact doubles); the rest takes the general path.  Returns 0 when the
// general path has to decide.
__device__ __forceinline__ int eat_plain_decimal(const Text &t, long long &q, long long lim, double *out)
{
    long long r = q;
    const bool neg = r < lim && t(r) == '-';
    if (neg) r++;
    const long long d0 = r;
    unsigned long long w = 0;
    int nsig = 0, frac = 0;
    while (r < lim) {
        const int c = t(r);
        if (c < '0' || c > '9') break;
        if (nsig > 0 || c != '0') nsig++;
        w = w * 10ull + (unsigned long long)(c - '0');
        r++;
        if (nsig > 15) return 0;
    }
    const int ni = (int)(r - d0);
    if (ni == 0 || (ni > 1 && t(d0) == '0')) return 0;
    if (r < lim && t(r) == '.') {
        r++;
        const long long f0 = r;
        while (r < lim) {
            const int c = t(r);
            if (c < '0' || c > '9') break;
            if (nsig > 0 || c != '0') nsig++;
            w = w * 10ull + (unsigned long long)(c - '0');
            r++;
            if (nsig > 15 || r - f0 > 22) return 0;
        }
        frac = (int)(r - f0);
        if (frac == 0) return 0;
    }
    if (r < lim && (t(r) == 'e' || t(r) == 'E')) return 0;
    const double mag = frac ? __ddiv_rn((double)w, c_json_pow10[frac]) : (double)w;
    *out = neg ? -mag : mag;
    q = r;
    return 1;
}

}  // namespace

constexpr int kParseStage = kThreads * kMaxChars;   // what 256 canonical texts can need

__global__ void __launch_bounds__(256)
view_from_json_kernel(const JsonReadParams p)
{
    __shared__ __align__(16) char stage[kParseStage + 16];
    const long long v0 = (long long)blockIdx.x * blockDim.x, v = v0 + threadIdx.x;
    // the CTA's texts are one contiguous run of characters: copy it in with coalesced loads
    const long long v1 = v0 + blockDim.x < p.n ? v0 + blockDim.x : p.n;
    const long long c_lo = p.offsets[v0], c_hi = p.offsets[v1];
    const long long a_lo = c_lo & ~15ll;                 // aligned start; never before the buffer's own start
    const bool staged = p.vec_ok && c_hi - a_lo <= kParseStage;
    if (staged) {
        for (long long o = a_lo + 16ll * threadIdx.x; o < c_hi; o += 16ll * blockDim.x) {
            uint4 w;
            if (o + 16 <= c_hi) {
                w = *reinterpret_cast<const uint4 *>(p.chars + o);
            } else {
                char tmp[16];
                for (int k = 0; k < 16; k++) tmp[k] = o + k < c_hi ? p.chars[o + k] : 0;
                memcpy(&w, tmp, 16);
            }
            *reinterpret_cast<uint4 *>(stage + (o - a_lo)) = w;
        }
    }
    __syncthreads();
    if (v >= p.n) return;
    const Text t{p.chars, staged ? stage : nullptr, a_lo};
    long long q = p.offsets[v];
    const long long lim = p.offsets[v + 1];
    int sn = 0, level = 0, url_id = 0, st = HLSMAP_JSON_OK;
    unsigned flags = 0;
    double time = __longlong_as_double(0x7ff8000000000000ll);
    bool first = true;
    do {
        if (!eat(t, q, lim, "{", 1)) {
            st = HLSMAP_JSON_MALFORMED;
            break;
        }
        if (eat(t, q, lim, "\"sn\":", 5)) {
            if ((st = eat_int(t, q, lim, &sn))) break;
            flags |= HLSMAP_JSON_HAS_SN;
            first = false;
        }
        if (eat(t, q, lim, first ? "\"trackView\":{" : ",\"trackView\":{", first ? 13 : 14)) {
            flags |= HLSMAP_JSON_HAS_TRACKVIEW;
            first = false;
            bool inner_first = true;
            if (eat(t, q, lim, "\"level\":", 8)) {
                if ((st = eat_int(t, q, lim, &level))) break;
                flags |= HLSMAP_JSON_HAS_LEVEL;
                inner_first = false;
            }
            if (eat(t, q, lim, inner_first ? "\"urlId\":" : ",\"urlId\":", inner_first ? 8 : 9)) {
                if ((st = eat_int(t, q, lim, &url_id))) break;
                flags |= HLSMAP_JSON_HAS_URLID;
            }
            if (!eat(t, q, lim, "}", 1)) {
                st = HLSMAP_JSON_MALFORMED;
                break;
            }
        }
        if (eat(t, q, lim, first ? "\"time\":" : ",\"time\":", first ? 7 : 8)) {
            flags |= HLSMAP_JSON_HAS_TIME;
            if (eat(t, q, lim, "null", 4)) {
                flags |= HLSMAP_JSON_TIME_NULL;
            } else {
                if (!eat_plain_decimal(t, q, lim, &time)) {
                    const JsonNumber n = scan_json_number(t, q, lim);
                    if (n.length == 0) {
                        st = HLSMAP_JSON_MALFORMED;
                        break;
                    }
                    q += n.length;
                    if (!json_number_to_double(n, &time)) {
                        st = HLSMAP_JSON_UNSUPPORTED;
                        break;
                    }
                }
            }
        }
        if (!eat(t, q, lim, "}", 1) || q != lim) st = HLSMAP_JSON_MALFORMED;
    } while (false);
    p.sn[v] = sn;
    p.level[v] = level;
    p.url_id[v] = url_id;
    p.time[v] = time;
    p.flags[v] = (unsigned char)flags;
    p.status[v] = (unsigned char)st;
}

int view_to_json_max_grid()
{
    static int grid = 0;
    if (!grid) {
        int dev = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, view_to_json_kernel, kThreads, 0) != cudaSuccess ||
            per_sm < 1)
            per_sm = 1;
        grid = per_sm * (sms > 0 ? sms : 148);
    }
    return grid;
}

cudaError_t launch_view_to_json(const JsonWriteParams &p, int grid, cudaStream_t s)
{
    view_to_json_kernel<<<grid, kThreads, 0, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_view_from_json(const JsonReadParams &p, cudaStream_t s)
{
    if (p.n <= 0) return cudaSuccess;
    view_from_json_kernel<<<(unsigned)((p.n + 255) / 256), 256, 0, s>>>(p);
    return cudaGetLastError();
}

}  // namespace hlsmap

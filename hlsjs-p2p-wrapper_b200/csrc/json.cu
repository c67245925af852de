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
#include "swar.cuh"
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

// Three launches, none of which waits for another CTA: (1) measure -- the digits of `time` and the
// length of every view's text, 12 bytes a view of staging and one sum per 256-view tile; (2) one
// CTA scans the tile sums; (3) render -- every tile knows where its characters start.  (An earlier
// single-pass build resolved the offsets with a decoupled look-back inside the kernel that also
// generates the digits; a third of its time went to waiting for the slowest tile ahead.)
__global__ void __launch_bounds__(kThreads)
view_json_measure_kernel(const JsonWriteParams p)
{
    __shared__ unsigned warp_sum[kThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long v = (long long)blockIdx.x * kThreads + tid;
    const bool in = v < p.n;
    int len = 0;
    if (in) {
        // time: 0 = property absent, 1 = null (NaN, +-Infinity), 2 = a number
        int tkind = 0, tlen = 0;
        bool tneg = false;
        Shortest sd;
        sd.digits = 0;
        sd.exp10 = 0;
        sd.ndigits = 1;
        if (p.time != nullptr) {
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
        len = 6 + int_len(p.sn[v]) + 22 + int_len(p.level[v]) + 9 + int_len(p.url_id[v]) + 1 + (tkind ? 8 + tlen : 0) + 1;
        p.stage_digits[v] = sd.digits;
        p.stage_meta[v] = ((unsigned)sd.exp10 & 0xffffu) | ((unsigned)sd.ndigits << 16) | ((unsigned)tkind << 21) |
                          ((unsigned)tneg << 23) | ((unsigned)len << 24);
    }
    int x = len;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) x += __shfl_xor_sync(0xffffffffu, x, d);
    if (lane == 0) warp_sum[warp] = (unsigned)x;
    __syncthreads();
    if (tid == 0) {
        unsigned t = 0;
#pragma unroll
        for (int w = 0; w < kThreads / 32; w++) t += warp_sum[w];
        p.block_sum[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(1024)
view_json_scan_kernel(const JsonWriteParams p)
{
    __shared__ unsigned long long wsum[32];
    __shared__ unsigned long long carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int b0 = 0; b0 < p.n_tiles; b0 += 1024) {
        const int b = b0 + tid;
        const unsigned long long v = b < p.n_tiles ? p.block_sum[b] : 0ull;
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
        if (b < p.n_tiles) p.block_base[b] = before + x - v;
        __syncthreads();
        if (tid == 0) carry += all;
        __syncthreads();
    }
    if (tid == 0) {
        const unsigned long long total = carry;
        p.offsets[p.n] = (int)total;
        p.result->total_hits = (long long)total;
        p.result->n_unparsed = 0;
        p.result->n_missing = 0;
        p.result->overflow = total > 0x7fffffffull ? 1 : 0;
    }
}

__global__ void __launch_bounds__(kThreads)
view_json_render_kernel(const JsonWriteParams p)
{
    __shared__ unsigned warp_cnt[kThreads / 32];
    __shared__ __align__(16) char buf[kThreads * kMaxChars + 16];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long v = (long long)blockIdx.x * kThreads + tid;
    const bool in = v < p.n;
    const unsigned meta = in ? p.stage_meta[v] : 0u;
    const int len = (int)(meta >> 24);
    int x = len;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
    }
    if (lane == 31) warp_cnt[warp] = (unsigned)x;
    __syncthreads();
    unsigned wbase = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; w++) {
        const unsigned c = warp_cnt[w];
        wbase += (w < warp) ? c : 0u;
        tile_total += c;
    }
    const unsigned long long tile_base = p.block_base[blockIdx.x];
    const int at = (int)wbase + x - len;
    // the tile's characters sit in shared memory at the same offset modulo 16 as their place in
    // global memory, so the copy out below can move 16 bytes per thread with aligned accesses
    const int pad = (int)((reinterpret_cast<unsigned long long>(p.chars) + tile_base) & 15ull);
    if (in) {
        p.offsets[v] = (int)(tile_base + (unsigned long long)at);
        const int tkind = (int)((meta >> 21) & 3u);
        char *d = buf + pad + at;
        d += put_lit(d, "{\"sn\":", 6);
        d += put_int(d, p.sn[v]);
        d += put_lit(d, ",\"trackView\":{\"level\":", 22);
        d += put_int(d, p.level[v]);
        d += put_lit(d, ",\"urlId\":", 9);
        d += put_int(d, p.url_id[v]);
        *d++ = '}';
        if (tkind) {
            d += put_lit(d, ",\"time\":", 8);
            Shortest sd;
            sd.digits = p.stage_digits[v];
            sd.exp10 = (int)(short)(meta & 0xffffu);
            sd.ndigits = (int)((meta >> 16) & 31u);
            if (tkind == 1) {
                d += put_lit(d, "null", 4);
            } else if (sd.digits == 0) {
                *d++ = '0';
            } else {
                if ((meta >> 23) & 1u) *d++ = '-';
                d += js_number_write(sd, d);
            }
        }
        *d++ = '}';
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

__constant__ unsigned long long c_json_ipow10[20] = {
    1ull, 10ull, 100ull, 1000ull, 10000ull, 100000ull, 1000000ull, 10000000ull, 100000000ull, 1000000000ull,
    10000000000ull, 100000000000ull, 1000000000000ull, 10000000000000ull, 100000000000000ull,
    1000000000000000ull, 10000000000000000ull, 100000000000000000ull, 1000000000000000000ull,
    10000000000000000000ull};

// `time`: digits [. digits] without an exponent and with at most 19 significant digits is read in
// one walk: up to 15 digits it is one exact IEEE division (both operands are exact doubles), up to
// 19 the exact integer division of decimal.cuh.  Returns 0 when the general path has to decide.
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
        if (nsig > 19) return 0;
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
            if (nsig > 19 || r - f0 > 22) return 0;
        }
        frac = (int)(r - f0);
        if (frac == 0) return 0;
    }
    if (r < lim && (t(r) == 'e' || t(r) == 'E')) return 0;
    double mag;
    if (w == 0) mag = 0.0;
    else if (nsig <= 15) mag = frac ? __ddiv_rn((double)w, c_json_pow10[frac]) : (double)w;
    else if (frac <= 19) mag = div_pow10_rounded(w, frac, false);
    else return 0;
    *out = neg ? -mag : mag;
    q = r;
    return 1;
}

// ---- the canonical text, eight bytes at a time ---------------------------------------------
// {"sn":<int>,"trackView":{"level":<int>,"urlId":<int>},"time":<number or null>} is what
// JSON.stringify gives for a SegmentView the constructor built (segment-view.js:22-26).  Its four
// literals are compared as 64-bit words and its three int32 values read with the SWAR routines of
// swar.cuh; anything that is not exactly this shape is left to the general reader below, which
// decides what it is.  Only for texts staged in shared memory.
struct Stage8 {
    const char *s;      // shared copy; s[0] is the character at absolute offset lo (16-byte aligned)
    long long lo;
    __device__ __forceinline__ unsigned long long operator()(long long p) const
    {
        const int r = (int)(p - lo);
        const unsigned *w = reinterpret_cast<const unsigned *>(s + (r & ~3));
        const unsigned a = w[0], b = w[1], c = w[2];
        const unsigned sh = (unsigned)(r & 3) * 8u;
        return (unsigned long long)__funnelshift_r(a, b, sh) | ((unsigned long long)__funnelshift_r(b, c, sh) << 32);
    }
};

__host__ __device__ constexpr unsigned long long lit8(const char *c, int n)
{
    unsigned long long v = 0;
    for (int i = 0; i < n && i < 8; i++) v |= (unsigned long long)(unsigned char)c[i] << (8 * i);
    return v;
}

// the literal `c` (n <= 24 characters) at q?  Advances q.
template <int N>
__device__ __forceinline__ bool eat8(const Stage8 &t, long long &q, long long lim, const char (&c)[N])
{
    constexpr int n = N - 1;
    if (q + n > lim) return false;
    bool ok = true;
#pragma unroll
    for (int o = 0; o < n; o += 8) {
        const int m = n - o < 8 ? n - o : 8;
        const unsigned long long mask = m == 8 ? ~0ull : ((1ull << (8 * m)) - 1ull);
        ok = ok && ((t(q + o) & mask) == lit8(c + o, m));
    }
    if (ok) q += n;
    return ok;
}

// The general reader: any subset of the properties in constructor order, every JSON number form.
// Kept out of line so that the canonical fast path of the kernel does not carry its registers.
__device__ __noinline__ void parse_view_general(const Text &t, long long q, long long lim, int &sn, int &level,
                                                int &url_id, double &time, unsigned &flags, int &st)
{
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
    if (staged) {       // the canonical full text in a few word operations; otherwise fall through
        const Stage8 t8{stage, a_lo};
        long long r = q;
        int a = 0, b = 0, c = 0;
        if (eat8(t8, r, lim, "{\"sn\":") && swar_int32(t8, r, lim, &a) &&
            eat8(t8, r, lim, ",\"trackView\":{\"level\":") && swar_int32(t8, r, lim, &b) &&
            eat8(t8, r, lim, ",\"urlId\":") && swar_int32(t8, r, lim, &c) && eat8(t8, r, lim, "},\"time\":")) {
            double x = time;
            unsigned f = HLSMAP_JSON_HAS_SN | HLSMAP_JSON_HAS_TRACKVIEW | HLSMAP_JSON_HAS_LEVEL |
                         HLSMAP_JSON_HAS_URLID | HLSMAP_JSON_HAS_TIME;
            bool ok = true;
            if (eat8(t8, r, lim, "null")) f |= HLSMAP_JSON_TIME_NULL;
            else ok = swar_plain_decimal(t8, r, lim, &x, c_json_pow10, c_json_ipow10) || eat_plain_decimal(t, r, lim, &x) != 0;
            if (ok && eat8(t8, r, lim, "}") && r == lim) {
                p.sn[v] = a;
                p.level[v] = b;
                p.url_id[v] = c;
                p.time[v] = x;
                p.flags[v] = (unsigned char)f;
                p.status[v] = (unsigned char)HLSMAP_JSON_OK;
                return;
            }
        }
    }
    parse_view_general(t, q, lim, sn, level, url_id, time, flags, st);
    p.sn[v] = sn;
    p.level[v] = level;
    p.url_id[v] = url_id;
    p.time[v] = time;
    p.flags[v] = (unsigned char)flags;
    p.status[v] = (unsigned char)st;
}

int view_to_json_max_grid() { return 0x7fffffff; }   // one CTA per 256-view tile: no persistent grid any more

cudaError_t launch_view_to_json(const JsonWriteParams &p, int, cudaStream_t s)
{
    if (p.n_tiles < 1) return cudaSuccess;
    view_json_measure_kernel<<<p.n_tiles, kThreads, 0, s>>>(p);
    view_json_scan_kernel<<<1, 1024, 0, s>>>(p);
    view_json_render_kernel<<<p.n_tiles, kThreads, 0, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_view_from_json(const JsonReadParams &p, cudaStream_t s)
{
    if (p.n <= 0) return cudaSuccess;
    view_from_json_kernel<<<(unsigned)((p.n + 255) / 256), 256, 0, s>>>(p);
    return cudaGetLastError();
}

}  // namespace hlsmap

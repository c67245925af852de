// ingest.cu -- media playlists (M3U8 text) -> fragment tables, on the device.
//
// the step before the MediaMap path.  In the reference deployment hls.js
// (package.json:34, "hls.js": "0.5.46", not vendored in the reference checkout) parses every
// level playlist into `level.details.fragments`, the objects media-map.js:28,41-48 reads.  Of
// that parser the mapping layer depends on exactly this much, restated from the published
// source of that release (src/loader/playlist-loader.js, parseLevelPlaylist):
//
//   one global regular expression walks the text left to right; of its alternatives
//     #EXT-X-MEDIA-SEQUENCE:(\d+)       currentSN = parseInt(digits)
//     #EXTINF:([\d.]+)[^\r\n]*...       duration = parseFloat(run); unless NaN, push
//                                       {sn: currentSN++, start: totalduration, duration}
//                                       and totalduration += duration
//     #EXT-X-ENDLIST                    level.live = false
//     #EXT-X-KEY:(.*)  #EXT-X-PROGRAM-DATE-TIME:(.*)     swallow the rest of their line
//   matter here; a match that swallows the rest of its line hides any tag later on that line.
//
// Work split.  m3u8_scan_kernel is byte-parallel and single pass: 8 KB tiles handed out by
// ticket, every '#' of a tile becomes a candidate, candidates are classified with all lanes
// busy, and two decoupled look-backs (fragments, sequence tags) give every fragment its global
// slot, so durations land in playlist order with the text read once.  What is inherently serial
// -- `start` is a running fp64 sum, which must be added in the reference's order to give the
// reference's bits -- is left to m3u8_assemble_kernel: one thread (short tracks) or one warp
// (long tracks) per track walks its durations once.
#include "common.cuh"
#include "decimal.cuh"
#include "scan.cuh"

namespace hlsmap {
namespace {

constexpr int kThreads = 256;
constexpr int kTile = kIngestTileBytes;       // bytes of text per tile
constexpr int kBytesPerThread = kTile / kThreads;  // 32
constexpr int kCap = 2048;                    // candidates classified per round
constexpr int kPerThread = kCap / kThreads;   // 8
constexpr int kWarps = kThreads / 32;
using Scan = TileScan<kThreads>;

enum TokKind { kTokNone = 0, kTokFrag = 1, kTokSeq = 2, kTokEnd = 3, kTokLine = 4, kTokInfNaN = 5 };

// Bytes of the text by absolute position: the tile's shared-memory copy when the position is
// inside it, global memory otherwise (tags running past the tile's end, backward line scans).
struct TextView {
    const unsigned char *g;
    const unsigned char *s;
    long long lo, hi;
    __device__ __forceinline__ int operator()(long long p) const
    {
        return (p >= lo && p < hi) ? (int)s[p - lo] : (int)g[p];
    }
};

__device__ __forceinline__ bool match_lit(const TextView &v, long long p, long long lim, const char *lit,
                                          int n)
{
    if (p + n > lim) return false;
    for (int i = 0; i < n; i++)
        if (v(p + i) != (int)(unsigned char)lit[i]) return false;
    return true;
}

struct Token {
    int kind;
    long long bits;   // kTokFrag: the duration's bit pattern; kTokSeq: the sequence number
    bool unsupported;
};

// What the regular expression matches at a '#' (position p of a text ending at lim).
__device__ __noinline__ Token classify_at(const TextView &v, long long p, long long lim)
{
    Token t;
    t.kind = kTokNone;
    t.bits = 0;
    t.unsupported = false;
    if (!match_lit(v, p, lim, "#EXT", 4)) return t;
    if (match_lit(v, p + 4, lim, "INF:", 4)) {
        const DecimalScan d = scan_decimal(v, p + 8, lim);
        if (d.run == 0) return t;          // [\d.]+ needs one character: no match at all
        if (!d.is_number) {                // parseFloat(".") is NaN: matched, no fragment
            t.kind = kTokInfNaN;
            return t;
        }
        double x = 0.0;
        t.unsupported = !decimal_to_double(d, &x);
        t.kind = kTokFrag;
        t.bits = __double_as_longlong(x);
        return t;
    }
    if (!match_lit(v, p + 4, lim, "-X-", 3)) return t;
    const long long q = p + 7;
    if (match_lit(v, q, lim, "MEDIA-SEQUENCE:", 15)) {
        long long val = 0;
        int nd = 0;
        for (long long r = q + 15; r < lim; r++) {
            const int c = v(r);
            if (c < '0' || c > '9') break;
            nd++;
            if (val <= 0x7fffffffll) val = val * 10 + (c - '0');   // saturates just past int32
        }
        if (nd == 0) return t;
        t.kind = kTokSeq;
        t.unsupported = val > 0x7fffffffll;
        t.bits = val > 0x7fffffffll ? 0x7fffffffll : val;
        return t;
    }
    if (match_lit(v, q, lim, "ENDLIST", 7)) {
        t.kind = kTokEnd;
        return t;
    }
    if (match_lit(v, q, lim, "KEY:", 4) || match_lit(v, q, lim, "PROGRAM-DATE-TIME:", 18)) t.kind = kTokLine;
    return t;
}

// A '#' in the middle of a line is a match only if nothing before it on the line swallowed it:
// replay the line from its first byte (global memory; this is the rare path).
__device__ __noinline__ bool hidden_by_earlier_match(const unsigned char *text, long long track_lo, long long p,
                                        long long lim)
{
    long long b = p;
    while (b > track_lo && text[b - 1] != '\n' && text[b - 1] != '\r') b--;
    TextView v{text, text, 0, 0};
    long long q = b;
    while (q < p) {
        if (text[q] != '#') {
            q++;
            continue;
        }
        const Token t = classify_at(v, q, lim);
        if (t.kind == kTokFrag || t.kind == kTokInfNaN) return true;   // [^\r\n]* runs to the line end
        if (t.kind == kTokLine) {
            // (.*) stops at a JS line terminator: \n, \r, U+2028, U+2029 (E2 80 A8 / E2 80 A9)
            long long r = q + 1;
            while (r < lim) {
                const int c = text[r];
                if (c == '\n' || c == '\r') break;
                if (c == 0xE2 && r + 2 < lim && text[r + 1] == 0x80 && (text[r + 2] == 0xA8 || text[r + 2] == 0xA9))
                    break;
                r++;
            }
            if (r >= p) return true;
            q = r;
            continue;
        }
        q++;
    }
    return false;
}

__device__ __forceinline__ unsigned hash_mask4(unsigned w)
{
    const unsigned m = __vcmpeq4(w, 0x23232323u);   // 0xff per '#'
    return ((m >> 7) & 1u) | ((m >> 14) & 2u) | ((m >> 21) & 4u) | ((m >> 28) & 8u);
}

struct Shared {
    Scan::Smem scan;
    __align__(16) unsigned char text[kTile];
    unsigned short cpos[kCap];          // tile-relative position of each candidate of the round
    unsigned short cfpre[kCap];         // fragments of the tile before that candidate
    unsigned wtot[kPerThread * kWarps]; // per (slot, warp): fragments | tags << 16
    unsigned wpre[kPerThread * kWarps + 1];
    int hash_warp[kWarps + 1];
};

struct RoundRegs {
    long long bits[kPerThread];
    unsigned pre[kPerThread];   // exclusive (fragments | tags << 16) within the tile
    unsigned char kind[kPerThread];
};

// Builds the candidate list of one round (the '#' bytes of threads [t0, t1)), classifies it and
// scans the fragment / tag flags.  Returns the round's totals as fragments | tags << 16.
__device__ unsigned classify_round(const IngestParams &p, Shared &sh, RoundRegs &rr, int tile, long long lo,
                                   long long hi, unsigned my_mask, int my_first, int nc, unsigned running,
                                   bool effects)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // candidate list, in text order
    if (my_first >= 0) {
        unsigned m = my_mask;
        int at = my_first;
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            sh.cpos[at++] = (unsigned short)(tid * kBytesPerThread + b);
        }
    }
    __syncthreads();
    const TextView v{p.text, sh.text, lo, hi};
    const int t_lo = p.tile_tlo[tile], t_hi = p.tile_tlo[tile + 1];
#pragma unroll
    for (int k = 0; k < kPerThread; k++) {
        const int i = tid + k * kThreads;
        int kind = kTokNone;
        long long bits = 0;
        if (i < nc) {
            const long long pos = lo + sh.cpos[i];
            int a = t_lo, b = t_hi;               // last track whose text starts at or before pos
            while (a < b) {
                const int mid = (a + b + 1) >> 1;
                if (p.text_off[mid] <= pos) a = mid;
                else b = mid - 1;
            }
            const long long tb = p.text_off[a], te = p.text_off[a + 1];
            const bool parsed = p.state == nullptr || p.state[a] == HLSMAP_TRACK_PARSED;
            if (parsed && pos < te) {
                const Token t = classify_at(v, pos, te);
                bool live = t.kind != kTokNone;
                if (live && pos > tb) {
                    const int prev = v(pos - 1);
                    if (prev != '\n' && prev != '\r') live = !hidden_by_earlier_match(p.text, tb, pos, te);
                }
                if (live) {
                    kind = t.kind;
                    bits = t.bits;
                    if (effects && t.unsupported) atomicAdd(&p.out->n_unsupported, 1ull);
                    if (effects && kind == kTokEnd) p.live[a] = 0;
                }
            }
        }
        rr.kind[k] = (unsigned char)kind;
        rr.bits[k] = bits;
        const unsigned fb = __ballot_sync(0xffffffffu, kind == kTokFrag);
        const unsigned tb2 = __ballot_sync(0xffffffffu, kind == kTokSeq);
        const unsigned below = (1u << lane) - 1u;
        rr.pre[k] = (unsigned)__popc(fb & below) | ((unsigned)__popc(tb2 & below) << 16);
        if (lane == 0) sh.wtot[k * kWarps + warp] = (unsigned)__popc(fb) | ((unsigned)__popc(tb2) << 16);
    }
    __syncthreads();
    if (warp == 0) {   // exclusive scan of the 64 (slot, warp) totals, slot-major = text order
        const unsigned a = sh.wtot[2 * lane], b = sh.wtot[2 * lane + 1];
        unsigned x = a + b;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        sh.wpre[2 * lane] = x - a - b;
        sh.wpre[2 * lane + 1] = x - b;
        if (lane == 31) sh.wpre[kPerThread * kWarps] = x;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kPerThread; k++) {
        const int i = tid + k * kThreads;
        rr.pre[k] += sh.wpre[k * kWarps + warp] + running;
        if (i < nc) sh.cfpre[i] = (unsigned short)(rr.pre[k] & 0xffffu);
    }
    const unsigned total = sh.wpre[kPerThread * kWarps];
    __syncthreads();
    return total;
}

// Writes what one classified round found: durations and sequence tags to their global slots,
// and the fragment offset of every track whose text starts inside [b_lo, b_hi) of the tile.
__device__ void emit_round(const IngestParams &p, Shared &sh, const RoundRegs &rr, int tile, long long lo,
                           int nc, unsigned round_end, unsigned long long fbase, unsigned long long tbase,
                           long long b_lo, long long b_hi)
{
    const int tid = threadIdx.x;
#pragma unroll
    for (int k = 0; k < kPerThread; k++) {
        const int i = tid + k * kThreads;
        if (i >= nc) continue;
        const unsigned long long f = fbase + (rr.pre[k] & 0xffffu);
        if (rr.kind[k] == kTokFrag) {
            if ((long long)f < p.capacity) p.dur[f] = __longlong_as_double(rr.bits[k]);
        } else if (rr.kind[k] == kTokSeq) {
            const unsigned long long j = tbase + (rr.pre[k] >> 16);
            if ((long long)j < p.tag_capacity) {
                p.tag_pos[j] = lo + sh.cpos[i];
                p.tag_frag[j] = (long long)f;
                p.tag_val[j] = (int)rr.bits[k];
            }
        }
    }
    // tracks starting in this byte range: fragments before the first candidate at/after the start
    const int u0 = p.tile_bfirst[tile], u1 = p.tile_bfirst[tile + 1];
    for (int u = u0 + tid; u < u1; u += kThreads) {
        const long long b = p.text_off[u];
        if (b < b_lo || b >= b_hi) continue;
        const long long rel = b - lo;
        int a = 0, z = nc;                       // first candidate with cpos >= rel
        while (a < z) {
            const int mid = (a + z) >> 1;
            if ((long long)sh.cpos[mid] < rel) a = mid + 1;
            else z = mid;
        }
        const unsigned fp = a < nc ? (unsigned)sh.cfpre[a] : (round_end & 0xffffu);
        p.track_off[u] = (long long)(fbase + fp);
    }
}

}  // namespace

__global__ void __launch_bounds__(kThreads, 4)
m3u8_scan_kernel(const IngestParams p)
{
    __shared__ Shared sh;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (;;) {
        const int tile = Scan::next_tile(sh.scan, p.ctr, p.tslot);
        if (tile >= p.n_tiles) break;
        const long long lo = (long long)tile * kTile;
        const long long hi = lo + kTile < p.n_bytes ? lo + kTile : p.n_bytes;
        const bool last = tile == p.n_tiles - 1;

        // stage the tile and find its '#' bytes
        unsigned mask = 0;
        {
            const long long at = lo + (long long)tid * kBytesPerThread;
#pragma unroll
            for (int g = 0; g < kBytesPerThread / 16; g++) {
                const long long a = at + g * 16;
                uint4 w = make_uint4(0, 0, 0, 0);
                if (p.vec_ok && a + 16 <= p.n_bytes) {
                    w = *reinterpret_cast<const uint4 *>(p.text + a);
                } else {
                    unsigned char tmp[16];
#pragma unroll
                    for (int j = 0; j < 16; j++) tmp[j] = a + j < p.n_bytes ? p.text[a + j] : 0;
                    memcpy(&w, tmp, 16);
                }
                *reinterpret_cast<uint4 *>(sh.text + tid * kBytesPerThread + g * 16) = w;
                mask |= (hash_mask4(w.x) | (hash_mask4(w.y) << 4) | (hash_mask4(w.z) << 8) |
                         (hash_mask4(w.w) << 12)) << (16 * g);
            }
        }
        // exclusive prefix of the per-thread '#' counts
        const int cnt = __popc(mask);
        int x = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) sh.hash_warp[warp] = x;
        __syncthreads();
        int before = x - cnt, n_hash = 0;
        int quarter_base[5];
#pragma unroll
        for (int w = 0; w < kWarps; w++) {
            const int c = sh.hash_warp[w];
            if ((w & 1) == 0) quarter_base[w >> 1] = n_hash;   // 2 warps = one quarter of the tile
            before += w < warp ? c : 0;
            n_hash += c;
        }
        quarter_base[4] = n_hash;

        RoundRegs rr;
        unsigned long long fbase = 0, tbase = 0;
        unsigned totals = 0;
        if (n_hash <= kCap) {
            // the usual case: one round, classified once, results kept in registers
            totals = classify_round(p, sh, rr, tile, lo, hi, mask, before, n_hash, 0u, true);
            Scan::publish_aggregate(p.status_f, p.epoch, tile, totals & 0xffffu);
            Scan::publish_aggregate(p.status_t, p.epoch, tile, totals >> 16);
            fbase = Scan::wait_prefix(sh.scan, p.status_f, p.epoch, tile, totals & 0xffffu);
            tbase = Scan::wait_prefix(sh.scan, p.status_t, p.epoch, tile, totals >> 16);
            emit_round(p, sh, rr, tile, lo, n_hash, totals, fbase, tbase, lo,
                       last ? 0x7fffffffffffffffll : hi);
        } else {
            // a tile dense with '#': four rounds of a quarter tile each, classified twice
            // (count, then write) so that no round holds more than kCap candidates
            for (int r = 0; r < 4; r++) {
                const bool mine = (warp >> 1) == r;
                totals += classify_round(p, sh, rr, tile, lo, hi, mine ? mask : 0u,
                                         mine ? before - quarter_base[r] : -1,
                                         quarter_base[r + 1] - quarter_base[r], 0u, false);
            }
            Scan::publish_aggregate(p.status_f, p.epoch, tile, totals & 0xffffu);
            Scan::publish_aggregate(p.status_t, p.epoch, tile, totals >> 16);
            fbase = Scan::wait_prefix(sh.scan, p.status_f, p.epoch, tile, totals & 0xffffu);
            tbase = Scan::wait_prefix(sh.scan, p.status_t, p.epoch, tile, totals >> 16);
            unsigned running = 0;
            for (int r = 0; r < 4; r++) {
                const bool mine = (warp >> 1) == r;
                const int nc = quarter_base[r + 1] - quarter_base[r];
                const unsigned t = classify_round(p, sh, rr, tile, lo, hi, mine ? mask : 0u,
                                                  mine ? before - quarter_base[r] : -1, nc, running, true);
                running += t;
                const long long b_lo = lo + (long long)r * (kTile / 4);
                const long long b_hi = (r == 3) ? (last ? 0x7fffffffffffffffll : hi) : b_lo + kTile / 4;
                emit_round(p, sh, rr, tile, lo, nc, running, fbase, tbase, b_lo, b_hi);
                __syncthreads();
            }
        }
        if (last && tid == 0) {
            p.out->n_fragments = fbase + (totals & 0xffffu);
            p.out->n_tags = tbase + (totals >> 16);
        }
    }
}

// First track of every tile, so that a candidate finds its track in a few steps.
__global__ void m3u8_tile_index_kernel(const long long *text_off, int n_tracks, int n_tiles, int *tile_tlo,
                                       int *tile_bfirst)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_tiles) return;
    if (i == n_tiles) {
        tile_tlo[i] = n_tracks > 0 ? n_tracks - 1 : 0;
        tile_bfirst[i] = n_tracks + 1;
        return;
    }
    const long long lo = (long long)i * kTile;
    int a = 0, b = n_tracks + 1;       // first index with text_off > lo
    while (a < b) {
        const int mid = (a + b) >> 1;
        if (text_off[mid] <= lo) a = mid + 1;
        else b = mid;
    }
    int t = a - 1;
    if (t > n_tracks - 1) t = n_tracks - 1;
    if (t < 0) t = 0;
    tile_tlo[i] = t;
    a = 0;
    b = n_tracks + 1;                  // first index with text_off >= lo
    while (a < b) {
        const int mid = (a + b) >> 1;
        if (text_off[mid] < lo) a = mid + 1;
        else b = mid;
    }
    tile_bfirst[i] = a;
}

// ---- per-track assembly: sn numbering and the running sum of durations --------------------
// LANES = 1: a thread per track (live windows of a few fragments); LANES = 32: a warp per track,
// coalesced loads, the 32 additions of a chunk replayed in order by every lane.
template <int LANES>
__global__ void __launch_bounds__(256)
m3u8_assemble_kernel(const AssembleParams p)
{
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long t = gid / LANES;
    const int lane = (int)(gid % LANES);
    if (t >= p.n_tracks) return;
    const bool parsed = p.state == nullptr || p.state[t] == HLSMAP_TRACK_PARSED;
    const long long off = p.track_off[t], len = p.track_off[t + 1] - off;
    if (lane == 0) p.dur0[t] = (parsed && len > 0) ? p.dur[off] : __longlong_as_double(0x7ff8000000000000ll);
    if (!parsed || len == 0) return;
    // sequence tags of this track: those whose position lies in its text
    const long long tb = p.text_off[t], te = p.text_off[t + 1];
    long long e = 0, z = p.n_tags;
    while (e < z) {
        const long long mid = (e + z) >> 1;
        if (p.tag_pos[mid] < tb) e = mid + 1;
        else z = mid;
    }
    long long cur = 0;
    double total = 0.0;
    bool overflow = false;
    for (long long c = 0; c < len; c += LANES) {
        const long long i = c + lane;
        const double d = i < len ? p.dur[off + i] : 0.0;
        const long long next_tag = (e < p.n_tags && p.tag_pos[e] < te) ? p.tag_frag[e] : 0x7fffffffffffffffll;
        const int n = len - c < LANES ? (int)(len - c) : LANES;
        long long my_sn = 0;
        double my_start = 0.0;
        if (next_tag > off + c + n - 1) {          // numbering runs straight through the chunk
            my_sn = cur + lane;
            cur += n;
            for (int j = 0; j < n; j++) {
                const double dj = LANES == 1 ? d : __shfl_sync(0xffffffffu, d, j);
                if (j == lane) my_start = total;
                total += dj;
            }
        } else {
            for (int j = 0; j < n; j++) {
                while (e < p.n_tags && p.tag_pos[e] < te && p.tag_frag[e] <= off + c + j) cur = p.tag_val[e++];
                const double dj = LANES == 1 ? d : __shfl_sync(0xffffffffu, d, j);
                if (j == lane) {
                    my_start = total;
                    my_sn = cur;
                }
                total += dj;
                cur++;
            }
        }
        if (i < len) {
            overflow |= my_sn > 0x7fffffffll;
            p.start[off + i] = my_start;
            p.sn[off + i] = (int)my_sn;
        }
    }
    if (overflow) atomicAdd(&p.out->n_unsupported, 1ull);
}

// ---- launchers ---------------------------------------------------------------------------
int ingest_scan_max_grid()
{
    static int grid = 0;
    if (!grid) {
        int dev = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, m3u8_scan_kernel, kThreads, 0) != cudaSuccess ||
            per_sm < 1)
            per_sm = 1;
        grid = per_sm * (sms > 0 ? sms : 148);
    }
    return grid;
}

cudaError_t launch_ingest_tile_index(const long long *text_off, int n_tracks, int n_tiles, int *tile_tlo,
                                     int *tile_bfirst, cudaStream_t s)
{
    const int n = n_tiles + 1;
    m3u8_tile_index_kernel<<<(n + 255) / 256, 256, 0, s>>>(text_off, n_tracks, n_tiles, tile_tlo, tile_bfirst);
    return cudaGetLastError();
}

cudaError_t launch_ingest_scan(const IngestParams &p, cudaStream_t s)
{
    int grid = ingest_scan_max_grid();
    if (grid > p.n_tiles) grid = p.n_tiles;
    if (grid < 1) return cudaSuccess;
    m3u8_scan_kernel<<<grid, kThreads, 0, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_ingest_assemble(const AssembleParams &p, bool warp_per_track, cudaStream_t s)
{
    if (p.n_tracks <= 0) return cudaSuccess;
    if (warp_per_track) {
        const long long threads = (long long)p.n_tracks * 32;
        m3u8_assemble_kernel<32><<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(p);
    } else {
        m3u8_assemble_kernel<1><<<(unsigned)((p.n_tracks + 255) / 256), 256, 0, s>>>(p);
    }
    return cudaGetLastError();
}

}  // namespace hlsmap

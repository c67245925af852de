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
// Work split.  m3u8_scan_kernel is byte-parallel and knows no dependency between tiles: 16 KB
// tiles, a 4 KB span per warp, every '#' of a span becomes a candidate, candidates are classified
// with all lanes busy, and each fragment's duration and each sequence tag is left in the span's
// fixed-size staging slot, with two counts per span.  A two-level scan of the counts and a
// warp-per-span copy put them at their final places (m3u8_block_scan / span_base / copy /
// track_offsets kernels), so the text is read once and nothing in the heavy kernel ever waits
// for another CTA.  What is inherently serial -- `start` is a running fp64 sum, which must be
// added in the reference's order to give the reference's bits -- is left to the assembly: a lane
// per track (m3u8_assemble_short_kernel for short tracks, m3u8_assemble_tracks_kernel for long ones;
// m3u8_assemble_kernel is the plain form both replaced, kept behind HLSMAP_ASSEMBLE_WARP).
#include <cstdlib>

#include "common.cuh"
#include "decimal.cuh"

namespace hlsmap {
namespace {

constexpr int kThreads = kIngestThreads;          // 128: four warps
constexpr int kTile = kIngestTileBytes;           // bytes of text per tile
constexpr int kWarps = kThreads / 32;
constexpr int kWarpBytes = kTile / kWarps;        // 4096: the span a warp owns
constexpr int kLaneBytes = kWarpBytes / 32;       // 128
constexpr int kGroups = kLaneBytes / 32;          // '#' masks per lane, one per 32 bytes
constexpr int kWarpCap = 256;                     // candidates a warp classifies per round
constexpr int kRoundBytes = kWarpCap;             // dense spans: text bytes per round
constexpr int kRounds = kWarpBytes / kRoundBytes;
constexpr int kLanesPerRound = kRoundBytes / kLaneBytes;
constexpr int kTrackStage = 130;                  // text offsets staged per tile

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
        t.unsupported = !decimal_to_double<false>(d, &x);   // rare here: the compact division will do
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

// The same decision for the tags that produce something (EXTINF, MEDIA-SEQUENCE, ENDLIST), taken
// from registers: up to 36 bytes at the '#' are fetched from the tile's shared-memory copy as
// aligned words and shifted into place, literals are compared a word at a time, numbers are
// walked with compile-time byte positions.  `limit` (<= 36) bytes are valid; `limit_is_end` says
// the text ends there.  Returns false when the token does not resolve within those bytes (a number
// longer than the window): the caller then takes classify_at().
__constant__ double c_pow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                   1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

__device__ __forceinline__ unsigned lit4(char a, char b, char c, char d)
{
    return (unsigned)(unsigned char)a | ((unsigned)(unsigned char)b << 8) | ((unsigned)(unsigned char)c << 16) |
           ((unsigned)(unsigned char)d << 24);
}

__device__ __noinline__ bool decimal_slow(unsigned long long w, int exp10, bool sticky, double *out)
{
    DecimalScan d;
    d.w = w;
    d.exp10 = exp10;
    d.sticky = sticky;
    d.run = 1;
    d.is_number = true;
    return decimal_to_double<false>(d, out);
}

__device__ __forceinline__ bool classify_fast(const unsigned char *stext, int pos_rel, int limit, bool limit_is_end,
                                              Token &t)
{
    t.kind = kTokNone;
    t.bits = 0;
    t.unsupported = false;
    if (limit < 9) return limit_is_end || limit >= 4;   // nothing that emits fits; "#EXT" test below needs 4
    const unsigned *w32 = reinterpret_cast<const unsigned *>(stext + (pos_rel & ~3));
    const unsigned sh = (unsigned)(pos_rel & 3) * 8u;
    unsigned a[10];
#pragma unroll
    for (int j = 0; j < 7; j++) a[j] = w32[j];
    const unsigned u0 = __funnelshift_r(a[0], a[1], sh), u1 = __funnelshift_r(a[1], a[2], sh);
    if (u0 != lit4('#', 'E', 'X', 'T')) return true;
    unsigned u[9];
#pragma unroll
    for (int j = 2; j < 6; j++) u[j] = __funnelshift_r(a[j], a[j + 1], sh);
    if (u1 == lit4('I', 'N', 'F', ':')) {
        // scan_decimal() on bytes 8..23, sixteen bytes at a time in two registers (decimal.cuh);
        // bytes past `limit` read as zero, and a window cut short is always cut by the text's end
        unsigned long long x0 = (unsigned long long)u[2] | ((unsigned long long)u[3] << 32);
        unsigned long long x1 = (unsigned long long)u[4] | ((unsigned long long)u[5] << 32);
        const int nv = limit - 8;
        if (nv < 8) {
            x0 &= (1ull << (8 * nv)) - 1ull;
            x1 = 0ull;
        } else if (nv < 16) {
            x1 &= nv == 8 ? 0ull : (1ull << (8 * (nv - 8))) - 1ull;
        }
        unsigned long long w = 0;
        int run = 0, nfrac = 0;
        bool is_number = false;
        const bool sticky = false;
        if (!swar_scan_decimal16(x0, x1, &run, &w, &nfrac, &is_number)) return false;
        const int exp10 = -nfrac;
        if (run == 0) return true;
        if (!is_number) {
            t.kind = kTokInfNaN;
            return true;
        }
        double x = 0.0;
        if (w == 0) x = 0.0;
        else if (w <= (1ull << 53)) x = __ddiv_rn((double)w, c_pow10[-exp10]);   // both exact: one IEEE division
        else t.unsupported = !decimal_slow(w, exp10, sticky, &x);
        t.kind = kTokFrag;
        t.bits = __double_as_longlong(x);
        return true;
    }
    if ((u1 & 0xffffffu) != (lit4('-', 'X', '-', 0) & 0xffffffu)) return true;
    const unsigned c7 = u1 >> 24;
    if (c7 == 'E') {
        if (limit >= 14 && u[2] == lit4('N', 'D', 'L', 'I') && (u[3] & 0xffffu) == (lit4('S', 'T', 0, 0) & 0xffffu))
            t.kind = kTokEnd;
        return true;
    }
    if (c7 == 'M') {
        if (limit < 23) return true;                    // "#EXT-X-MEDIA-SEQUENCE:" and one digit
        if (u[2] != lit4('E', 'D', 'I', 'A') || u[3] != lit4('-', 'S', 'E', 'Q') || u[4] != lit4('U', 'E', 'N', 'C') ||
            (u[5] & 0xffffu) != (lit4('E', ':', 0, 0) & 0xffffu))
            return true;
#pragma unroll
        for (int j = 7; j < 10; j++) a[j] = w32[j];
#pragma unroll
        for (int j = 6; j < 9; j++) u[j] = __funnelshift_r(a[j], a[j + 1], sh);
        long long val = 0;
        int nd = 0;
        bool ended = false;
#pragma unroll
        for (int j = 22; j < 36; j++) {
            if (j >= limit) {
                if (!limit_is_end) return false;
                ended = true;
                break;
            }
            const int c = (int)((u[j >> 2] >> ((j & 3) * 8)) & 0xffu);
            if (c < '0' || c > '9') {
                ended = true;
                break;
            }
            nd++;
            if (val <= 0x7fffffffll) val = val * 10 + (c - '0');
        }
        if (!ended) return false;
        if (nd == 0) return true;
        t.kind = kTokSeq;
        t.unsupported = val > 0x7fffffffll;
        t.bits = val > 0x7fffffffll ? 0x7fffffffll : val;
        return true;
    }
    return true;   // KEY / PROGRAM-DATE-TIME and every other tag emit nothing themselves
}

// A '#' in the middle of a line is a match only if nothing before it on the line swallowed it:
// replay the line from its first byte (global memory; this is the rare path).
// A line is replayed once per '#' inside it by ONE thread, so the work is quadratic in the line's
// length: playlist text comes from the network, and one multi-megabyte line packed with tags could
// stall the kernel for seconds.  No real playlist has lines anywhere near kMaxReplayLine bytes (a
// segment URI or a key URI is a few hundred); a '#' further than that from its line start is refused
// (`too_long`, counted in n_unsupported -> HLSMAP_ERR_UNSUPPORTED), never guessed.
constexpr long long kMaxReplayLine = 16384;
__device__ __noinline__ bool hidden_by_earlier_match(const unsigned char *text, long long track_lo, long long p,
                                        long long lim, bool &too_long)
{
    long long b = p;
    too_long = false;
    while (b > track_lo && text[b - 1] != '\n' && text[b - 1] != '\r') {
        b--;
        if (p - b > kMaxReplayLine) {
            too_long = true;
            return true;
        }
    }
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

// '#' bytes of 16 bytes of text as 16 bits in text order.  Per word: compare, keep one bit a byte,
// and gather the four bits with one multiply (the partial products never collide below bit 28).
__device__ __forceinline__ unsigned hash_nibble(unsigned w)
{
    return ((__vcmpeq4(w, 0x23232323u) & 0x01010101u) * 0x01020408u) >> 24;
}
__device__ __forceinline__ unsigned hash_mask16(const uint4 &w)
{
    return hash_nibble(w.x) | (hash_nibble(w.y) << 4) | (hash_nibble(w.z) << 8) | (hash_nibble(w.w) << 12);
}

// text_off[] of the tracks a tile touches, from shared memory when they were staged.
struct TrackOffsets {
    const long long *g;
    const long long *s;
    int lo, n;
    __device__ __forceinline__ long long operator()(int t) const
    {
        const unsigned d = (unsigned)(t - lo);
        return d < (unsigned)n ? s[d] : g[t];
    }
};

// Shared state.  A tile is cut into one 4 KB span per warp; after the one barrier that makes the
// whole tile's text visible (tokens may run from one span into the next) a warp works alone on its
// span, with its own lists, and nothing in this kernel waits for another tile.
struct WarpShared {
    unsigned short cpos[kWarpCap];   // tile-relative position of each candidate of the round
    unsigned short cfpre[kWarpCap];  // fragments of the span before that candidate
    unsigned short ctrk[kWarpCap];   // track of the owning lane's first byte, relative to the tile's first track
};
struct Shared {
    __align__(16) unsigned char text[kTile + 48];   // + room for the word fetches of classify_fast
    long long toff[kTrackStage];
    WarpShared w[kWarps];
};

// One candidate: its track, what the expression matches there, whether an earlier match on the
// line hides it.
__device__ __forceinline__ int classify_candidate(const IngestParams &p, const Shared &sh, const TrackOffsets &toff,
                                                  int t_lo, int t_hi, long long lo, long long hi, int pos_rel,
                                                  bool effects, long long &bits)
{
    const TextView v{p.text, sh.text, lo, hi};
    int kind = kTokNone;
    bits = 0;
    const long long pos = lo + pos_rel;
    int a = t_lo;                         // t_lo: a track at or before the one holding pos (the owner lane's
    while (a < t_hi && toff(a + 1) <= pos) a++;   // hint); the last track starting at or before pos is a few steps on
    const long long tb = toff(a), te = toff(a + 1);
    const bool parsed = p.state == nullptr || p.state[a] == HLSMAP_TRACK_PARSED;
    if (parsed && pos < te) {
        // bytes valid both in the tile's copy and in the track; a candidate in the tile's last
        // bytes whose track runs on is left to the byte-wise classifier
        const long long in_tile = hi - pos, in_track = te - pos;
        Token t;
        bool done = false;
        if (in_track <= in_tile || in_tile >= 36) {
            const int limit = in_track < 36 ? (int)in_track : 36;
            done = classify_fast(sh.text, pos_rel, limit, limit == in_track, t);
        }
        if (!done) t = classify_at(v, pos, te);
        bool live = t.kind != kTokNone;
        if (live && pos > tb) {
            const int prev = v(pos - 1);
            if (prev != '\n' && prev != '\r') {
                bool too_long;
                live = !hidden_by_earlier_match(p.text, tb, pos, te, too_long);
                if (too_long && effects) atomicAdd(&p.out->n_unsupported, 1ull);
            }
        }
        if (live) {
            kind = t.kind;
            bits = t.bits;
            if (effects && t.unsupported) atomicAdd(&p.out->n_unsupported, 1ull);
            if (effects && kind == kTokEnd) p.live[a] = 0;
        }
    }
    return kind;
}

// Warp-wide: lists this warp's '#' bytes (those of the lanes with my_first >= 0) in text order,
// classifies them 32 at a time and leaves every fragment's duration and every sequence tag in
// the span's staging slot, at its rank within the span.  Returns the round's totals (fragments |
// tags << 16); `running` = the span's totals from earlier rounds.
__device__ __forceinline__ unsigned warp_classify(const IngestParams &p, Shared &sh, WarpShared &ws,
                                                  const TrackOffsets &toff, int t_lo, int t_hi, long long lo,
                                                  long long hi, const unsigned (&mask)[kGroups], int lane_base,
                                                  int my_first, int nc, unsigned running, long long span)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (my_first >= 0) {
        // the track holding this lane's first byte, searched once: its candidates start from there
        int a = t_lo, b2 = t_hi;
        const long long first_byte = lo + lane_base;
        while (a < b2) {
            const int mid = (a + b2 + 1) >> 1;
            if (toff(mid) <= first_byte) a = mid;
            else b2 = mid - 1;
        }
        const unsigned short hint = (unsigned short)(a - t_lo < 0xffff ? a - t_lo : 0xffff);
        int at = my_first;
#pragma unroll
        for (int g = 0; g < kGroups; g++) {
            unsigned m = mask[g];
            while (m) {
                const int b = __ffs(m) - 1;
                m &= m - 1;
                ws.ctrk[at] = hint;
                ws.cpos[at++] = (unsigned short)(lane_base + g * 32 + b);
            }
        }
    }
    __syncwarp();
    unsigned run = 0;
    const unsigned below = (1u << lane) - 1u;
    for (int i0 = 0; i0 < nc; i0 += 32) {
        const int i = i0 + lane;
        long long bits = 0;
        const int pos_rel = i < nc ? (int)ws.cpos[i] : 0;
        const int kind = i < nc ? classify_candidate(p, sh, toff, t_lo + (int)ws.ctrk[i < nc ? i : 0], t_hi, lo, hi,
                                                     pos_rel, true, bits)
                                : kTokNone;
        const unsigned fb = __ballot_sync(0xffffffffu, kind == kTokFrag);
        const unsigned tb = __ballot_sync(0xffffffffu, kind == kTokSeq);
        if (i < nc) {
            const unsigned pre = running + run + ((unsigned)__popc(fb & below) | ((unsigned)__popc(tb & below) << 16));
            const unsigned fpre = pre & 0xffffu, tpre = pre >> 16;
            ws.cfpre[i] = (unsigned short)fpre;
            if (kind == kTokFrag && p.stage_dur != nullptr && fpre < (unsigned)kIngestSpanFrags)
                p.stage_dur[span * kIngestSpanFrags + fpre] = __longlong_as_double(bits);
            if (kind == kTokSeq && p.stage_tag != nullptr && tpre < (unsigned)kIngestSpanTags)
                p.stage_tag[span * kIngestSpanTags + tpre] =
                    (unsigned long long)(unsigned)bits | ((unsigned long long)fpre << 32) |
                    ((unsigned long long)(unsigned)(pos_rel - warp * kWarpBytes) << 48);
        }
        run += (unsigned)__popc(fb) | ((unsigned)__popc(tb) << 16);
    }
    __syncwarp();
    return run;
}

// Warp-wide: the fragment offset, relative to the span, of every track whose text starts inside
// [b_lo, b_hi) -- the fragments before the first candidate at or after the track's first byte.
// m3u8_track_offsets_kernel adds the span's base afterwards.
__device__ __forceinline__ void warp_bounds(const IngestParams &p, const WarpShared &ws, const TrackOffsets &toff,
                                            int u0, int u1, long long lo, int nc, unsigned round_end, long long b_lo,
                                            long long b_hi)
{
    const int lane = threadIdx.x & 31;
    for (int u = u0 + lane; u < u1; u += 32) {
        const long long b = toff(u);
        if (b < b_lo || b >= b_hi) continue;
        const long long rel = b - lo;
        int a = 0, z = nc;                       // first candidate with cpos >= rel
        while (a < z) {
            const int mid = (a + z) >> 1;
            if ((long long)ws.cpos[mid] < rel) a = mid + 1;
            else z = mid;
        }
        p.track_off[u] = (long long)(a < nc ? (unsigned)ws.cfpre[a] : (round_end & 0xffffu));
    }
}

}  // namespace

__global__ void __launch_bounds__(kThreads, 8)
m3u8_scan_kernel(const IngestParams p)
{
    __shared__ Shared sh;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    WarpShared &ws = sh.w[warp];
    const long long open_end = 0x7fffffffffffffffll;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        const long long lo = (long long)tile * kTile;
        const long long hi = lo + kTile < p.n_bytes ? lo + kTile : p.n_bytes;
        const bool last = tile == p.n_tiles - 1;
        __syncthreads();     // the previous tile's text is no longer read by anyone

        // stage the tile (each warp its own 4 KB span, coalesced 16-byte pieces) and the text
        // offsets of the tracks the tile touches
#pragma unroll
        for (int j = 0; j < kWarpBytes / 512; j++) {
            const int piece = warp * (kWarpBytes / 16) + j * 32 + lane;
            const long long a = lo + (long long)piece * 16;
            uint4 w = make_uint4(0, 0, 0, 0);
            if (p.vec_ok && a + 16 <= p.n_bytes) {
                w = __ldcs(reinterpret_cast<const uint4 *>(p.text + a));
            } else if (a < p.n_bytes) {
                unsigned char tmp[16];
#pragma unroll
                for (int q = 0; q < 16; q++) tmp[q] = a + q < p.n_bytes ? p.text[a + q] : 0;
                memcpy(&w, tmp, 16);
            }
            *reinterpret_cast<uint4 *>(sh.text + piece * 16) = w;
        }
        {   // the tile this CTA takes next: one 128-byte line per thread on its way to L2
            const long long nx = lo + (long long)gridDim.x * kTile + (long long)tid * 128;
            if (nx < p.n_bytes) asm volatile("prefetch.global.L2 [%0];" ::"l"(p.text + nx));
        }
        const int t_lo = p.tile_tlo[tile], t_hi = p.tile_tlo[tile + 1];
        const int u0 = p.tile_bfirst[tile], u1 = p.tile_bfirst[tile + 1];
        const int n_staged = t_hi + 2 - t_lo < kTrackStage ? t_hi + 2 - t_lo : kTrackStage;
        for (int i = tid; i < n_staged; i += kThreads) sh.toff[i] = p.text_off[t_lo + i];
        const TrackOffsets toff{p.text_off, sh.toff, t_lo, n_staged};
        __syncthreads();     // tokens may run from one warp's span into the next

        // the '#' bytes of this lane's 128 bytes, and their rank within the warp's span
        const int lane_base = warp * kWarpBytes + lane * kLaneBytes;
        unsigned mask[kGroups];
        int cnt = 0;
#pragma unroll
        for (int g = 0; g < kGroups; g++) {
            const unsigned char *at = sh.text + lane_base + g * 32;
            mask[g] = hash_mask16(*reinterpret_cast<const uint4 *>(at)) |
                      (hash_mask16(*reinterpret_cast<const uint4 *>(at + 16)) << 16);
            cnt += __popc(mask[g]);
        }
        int x = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        const int before = x - cnt;
        const int n_hash = __shfl_sync(0xffffffffu, x, 31);
        const long long span = (long long)tile * kWarps + warp;
        const long long w_lo = lo + (long long)warp * kWarpBytes;
        const long long w_hi = warp == kWarps - 1 ? (last ? open_end : lo + kTile) : w_lo + kWarpBytes;

        unsigned totals = 0;
        if (n_hash <= kWarpCap) {
            totals = warp_classify(p, sh, ws, toff, t_lo, t_hi, lo, hi, mask, lane_base, before, n_hash, 0u, span);
            warp_bounds(p, ws, toff, u0, u1, lo, n_hash, totals, w_lo, w_hi);
        } else {
            // a span dense with '#': rounds of 256 bytes, never more than kWarpCap candidates each
            for (int r = 0; r < kRounds; r++) {
                const int r0 = __shfl_sync(0xffffffffu, before, r * kLanesPerRound);
                const int r1n = __shfl_sync(0xffffffffu, before, ((r + 1) * kLanesPerRound) & 31);
                const int r1 = r == kRounds - 1 ? n_hash : r1n;
                const bool mine = lane / kLanesPerRound == r;
                totals += warp_classify(p, sh, ws, toff, t_lo, t_hi, lo, hi, mask, lane_base, mine ? before - r0 : -1,
                                        r1 - r0, totals, span);
                const long long b_lo = w_lo + (long long)r * kRoundBytes;
                const long long b_hi = r == kRounds - 1 ? w_hi : b_lo + kRoundBytes;
                warp_bounds(p, ws, toff, u0, u1, lo, r1 - r0, totals, b_lo, b_hi);
                __syncwarp();
            }
        }
        if (lane == 0) {
            p.span_cnt[span] = totals;
            const long long blk = span / kIngestScanBlock;
            if (totals & 0xffffu) atomicAdd(&p.block_sum[2 * blk], (unsigned long long)(totals & 0xffffu));
            if (totals >> 16) atomicAdd(&p.block_sum[2 * blk + 1], (unsigned long long)(totals >> 16));
        }
    }
}

// ---- from per-span slots to the final columns ----------------------------------------------
// Two-level exclusive scan of the span counts: m3u8_scan_kernel already summed them per block of
// 1024 spans (atomics); one CTA scans the block sums, then every block rescans its own spans.
__global__ void __launch_bounds__(1024)
m3u8_block_scan_kernel(const IngestParams p)
{
    __shared__ unsigned long long wsum[2][32];
    __shared__ unsigned long long carry[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 2) carry[tid] = 0;
    __syncthreads();
    for (int b0 = 0; b0 < p.n_blocks; b0 += 1024) {
        const int b = b0 + tid;
        unsigned long long v[2], x[2];
#pragma unroll
        for (int a = 0; a < 2; a++) {
            v[a] = b < p.n_blocks ? p.block_sum[2 * b + a] : 0ull;
            x[a] = v[a];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned long long y = __shfl_up_sync(0xffffffffu, x[a], d);
                if (lane >= d) x[a] += y;
            }
            if (lane == 31) wsum[a][warp] = x[a];
        }
        __syncthreads();
#pragma unroll
        for (int a = 0; a < 2; a++) {
            unsigned long long before = carry[a], all = 0;
            for (int w = 0; w < 32; w++) {
                const unsigned long long c = wsum[a][w];
                before += w < warp ? c : 0ull;
                all += c;
            }
            if (b < p.n_blocks) p.block_base[2 * b + a] = before + x[a] - v[a];
            __syncthreads();
            if (tid == 0) carry[a] += all;
        }
        __syncthreads();
    }
    if (tid == 0) {
        p.block_base[2 * p.n_blocks] = carry[0];
        p.block_base[2 * p.n_blocks + 1] = carry[1];
        p.out->n_fragments = carry[0];
        p.out->n_tags = carry[1];
    }
}

// One CTA per block of 1024 spans: scan the block's span counts on top of the block's base.
__global__ void __launch_bounds__(256)
m3u8_span_base_kernel(const IngestParams p)
{
    __shared__ unsigned wsum[2][8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long s0 = (long long)blockIdx.x * kIngestScanBlock;
    unsigned c[4], f[4], t[4], fs = 0, ts = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const long long s = s0 + tid * 4 + k;
        c[k] = s < p.n_spans ? p.span_cnt[s] : 0u;
        f[k] = fs;
        t[k] = ts;
        fs += c[k] & 0xffffu;
        ts += c[k] >> 16;
    }
    unsigned xf = fs, xt = ts;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned yf = __shfl_up_sync(0xffffffffu, xf, d), yt = __shfl_up_sync(0xffffffffu, xt, d);
        if (lane >= d) {
            xf += yf;
            xt += yt;
        }
    }
    if (lane == 31) {
        wsum[0][warp] = xf;
        wsum[1][warp] = xt;
    }
    __syncthreads();
    unsigned bf = xf - fs, bt = xt - ts;
    for (int w = 0; w < warp; w++) {
        bf += wsum[0][w];
        bt += wsum[1][w];
    }
    const unsigned long long base_f = p.block_base[2 * blockIdx.x], base_t = p.block_base[2 * blockIdx.x + 1];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const long long s = s0 + tid * 4 + k;
        if (s < p.n_spans) {
            p.span_base[2 * s] = base_f + bf + f[k];
            p.span_base[2 * s + 1] = base_t + bt + t[k];
        }
    }
}

// A warp per span: the span's staged durations and tags go to their final, compact places
// (coalesced both ways).
__global__ void __launch_bounds__(256)
m3u8_copy_kernel(const IngestParams p)
{
    const int lane = threadIdx.x & 31;
    const long long s = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (s >= p.n_spans) return;
    const unsigned cnt = p.span_cnt[s];
    const int nf = (int)(cnt & 0xffffu), nt = (int)(cnt >> 16);
    const unsigned long long f0 = p.span_base[2 * s], t0 = p.span_base[2 * s + 1];
    for (int j = lane; j < nf; j += 32)
        if ((long long)(f0 + j) < p.capacity) p.dur[f0 + j] = p.stage_dur[s * kIngestSpanFrags + j];
    for (int j = lane; j < nt; j += 32) {
        if ((long long)(t0 + j) >= p.tag_capacity) continue;
        const unsigned long long e = p.stage_tag[s * kIngestSpanTags + j];
        p.tag_val[t0 + j] = (int)(unsigned)e;
        p.tag_frag[t0 + j] = (long long)(f0 + ((e >> 32) & 0xffffull));
        p.tag_pos[t0 + j] = s * kIngestSpanBytes + (long long)(e >> 48);
    }
}

// track_off[u] holds the fragments of u's span that come before u's first byte: add the span's base.
__global__ void m3u8_track_offsets_kernel(const IngestParams p)
{
    const long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (u > p.n_tracks) return;
    long long s = p.text_off[u] / kIngestSpanBytes;
    if (s > p.n_spans - 1) s = p.n_spans - 1;
    p.track_off[u] += (long long)p.span_base[2 * s];
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
    __shared__ __align__(16) double chunk[LANES == 1 ? 1 : 8][32];
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long t = gid / LANES;
    const int lane = (int)(gid % LANES);
    if (t >= p.n_tracks) return;
    // launched without a host round trip after the scan: step aside if the caller's columns or the
    // tag list turned out too small (the host sees the same counters and reports / retries)
    const long long n_tags_seen = (long long)p.out->n_tags;
    if ((long long)p.out->n_fragments > p.capacity || n_tags_seen > p.tag_capacity) return;
    const bool parsed = p.state == nullptr || p.state[t] == HLSMAP_TRACK_PARSED;
    const long long off = p.track_off[t], len = p.track_off[t + 1] - off;
    if (lane == 0) p.dur0[t] = (parsed && len > 0) ? p.dur[off] : __longlong_as_double(0x7ff8000000000000ll);
    if (!parsed || len == 0) return;
    // sequence tags of this track: those whose position lies in its text
    const long long tb = p.text_off[t], te = p.text_off[t + 1];
    // (first tag at or after the track's first byte; with one tag per playlist, as usual, that is
    // tag number t, which two independent loads confirm -- else the binary search)
    long long e = 0, z = n_tags_seen;
    if (t < n_tags_seen && p.tag_pos[t] >= tb && (t == 0 || p.tag_pos[t - 1] < tb)) {
        e = t;
    } else {
        while (e < z) {
            const long long mid = (e + z) >> 1;
            if (p.tag_pos[mid] < tb) e = mid + 1;
            else z = mid;
        }
    }
    long long cur = 0;
    double total = 0.0;
    bool overflow = false;
    const long long none = 0x7fffffffffffffffll;
    long long next_tag = (e < n_tags_seen && p.tag_pos[e] < te) ? p.tag_frag[e] : none;
    constexpr int kAhead = 4;                  // chunks whose loads are in flight together
    double ahead[kAhead];
#pragma unroll
    for (int k = 0; k < kAhead; k++) ahead[k] = lane + k * LANES < len ? p.dur[off + lane + k * LANES] : 0.0;
    for (long long c = 0; c < len; c += LANES) {
        const long long i = c + lane;
        const int slot = (int)((c / LANES) % kAhead);
        double d = 0.0;
#pragma unroll
        for (int k = 0; k < kAhead; k++) d = slot == k ? ahead[k] : d;
        if (slot == kAhead - 1) {              // the window is used up: fetch the next kAhead chunks at once
#pragma unroll
            for (int k = 0; k < kAhead; k++) {
                const long long ii = i + (long long)(k + 1) * LANES;
                ahead[k] = ii < len ? p.dur[off + ii] : 0.0;
            }
        }
        const int n = len - c < LANES ? (int)(len - c) : LANES;
        // tags sitting right at the chunk's first fragment (the usual one: the playlist's header)
        while (next_tag <= off + c) {
            cur = p.tag_val[e++];
            next_tag = (e < n_tags_seen && p.tag_pos[e] < te) ? p.tag_frag[e] : none;
        }
        long long my_sn = 0;
        double my_start = 0.0;
        if (next_tag > off + c + n - 1) {          // numbering runs straight through the chunk
            my_sn = cur + lane;
            cur += n;
            if (LANES == 1) {
                my_start = total;
                total += d;
            } else {
                // Lane L needs total + d0 + ... + d(L-1), added in that order.  The 32 durations go
                // through shared memory (one broadcast 16-byte load brings two of them to every
                // lane) and each lane adds only the ones before its own: 2.5 instructions per
                // fragment instead of shuffling every value to a chain every lane replays in full.
                // Lanes past the end hold 0.0, and x + 0.0 == x, so the loop has a fixed length.
                double *mine = chunk[threadIdx.x >> 5];
                __syncwarp();
                mine[lane] = d;
                __syncwarp();
                double acc = total;
#pragma unroll
                for (int j = 0; j < LANES; j += 2) {
                    const double2 dd = *reinterpret_cast<const double2 *>(mine + j);
                    if (j < lane) acc += dd.x;
                    if (j + 1 < lane) acc += dd.y;
                }
                my_start = acc;
                total = __shfl_sync(0xffffffffu, acc + d, LANES - 1);
            }
        } else {                                   // a tag in the middle of the chunk
            for (int j = 0; j < n; j++) {
                while (next_tag <= off + c + j) {
                    cur = p.tag_val[e++];
                    next_tag = (e < n_tags_seen && p.tag_pos[e] < te) ? p.tag_frag[e] : none;
                }
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

// ---- short tracks: a lane per track, 32 consecutive tracks per warp ---------------------------
// Consecutive tracks own consecutive fragments, so the fragments of a warp's 32 tracks are one
// contiguous run: when it is short enough (live windows: 6 fragments a track) it is loaded into
// shared memory with coalesced reads, each lane walks its own track there, and start / sn leave
// with coalesced writes.  A longer run is walked in global memory, a track per lane, as before.
constexpr int kShortWarps = 4, kShortCap = 512;
__global__ void __launch_bounds__(32 * kShortWarps)
m3u8_assemble_short_kernel(const AssembleParams p)
{
    __shared__ double s_val[kShortWarps][kShortCap];     // durations in, starts out
    __shared__ int s_sn[kShortWarps][kShortCap];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long t0 = ((long long)blockIdx.x * kShortWarps + warp) * 32;
    if (t0 >= p.n_tracks) return;
    const long long n_tags_seen = (long long)p.out->n_tags;
    if ((long long)p.out->n_fragments > p.capacity || n_tags_seen > p.tag_capacity) return;
    const long long t = t0 + lane;
    const bool valid = t < p.n_tracks;
    const bool parsed = valid && (p.state == nullptr || p.state[t] == HLSMAP_TRACK_PARSED);
    const long long off = valid ? p.track_off[t] : 0;
    const long long len = parsed ? p.track_off[t + 1] - off : 0;
    const long long t_end = t0 + 32 < p.n_tracks ? t0 + 32 : p.n_tracks;
    const long long run_lo = p.track_off[t0], run_n = p.track_off[t_end] - run_lo;
    const bool in_smem = run_n <= kShortCap;              // warp-uniform
    if (in_smem) {
        for (long long i = lane; i < run_n; i += 32) s_val[warp][i] = p.dur[run_lo + i];
        __syncwarp();
    }
    const double *src = in_smem ? &s_val[warp][off - run_lo] : p.dur + off;
    double *dst_start = in_smem ? &s_val[warp][off - run_lo] : p.start + off;
    int *dst_sn = in_smem ? &s_sn[warp][off - run_lo] : p.sn + off;
    if (valid) p.dur0[t] = len > 0 ? src[0] : __longlong_as_double(0x7ff8000000000000ll);
    if (len > 0) {
        // sequence tags of this track (see m3u8_assemble_kernel)
        const long long tb = p.text_off[t], te = p.text_off[t + 1];
        const long long none = 0x7fffffffffffffffll;
        long long e = 0, z = n_tags_seen;
        if (t < n_tags_seen && p.tag_pos[t] >= tb && (t == 0 || p.tag_pos[t - 1] < tb)) {
            e = t;
        } else {
            while (e < z) {
                const long long mid = (e + z) >> 1;
                if (p.tag_pos[mid] < tb) e = mid + 1;
                else z = mid;
            }
        }
        long long next_tag = (e < n_tags_seen && p.tag_pos[e] < te) ? p.tag_frag[e] : none;
        long long cur = 0;
        double total = 0.0;
        bool overflow = false;
        for (long long i = 0; i < len; i++) {
            while (next_tag <= off + i) {
                cur = p.tag_val[e++];
                next_tag = (e < n_tags_seen && p.tag_pos[e] < te) ? p.tag_frag[e] : none;
            }
            const double d = src[i];
            dst_start[i] = total;
            total += d;
            overflow |= cur > 0x7fffffffll;
            dst_sn[i] = (int)cur;
            cur++;
        }
        if (overflow) atomicAdd(&p.out->n_unsupported, 1ull);
    }
    if (in_smem) {
        __syncwarp();
        for (long long i = lane; i < run_n; i += 32) {
            p.start[run_lo + i] = s_val[warp][i];
            p.sn[run_lo + i] = s_sn[warp][i];
        }
    }
}

// ---- long tracks: 32 tracks per CTA, a lane per track ------------------------------------------
// The running sum of a track is a chain nobody can shorten, but 32 lanes can each walk their own
// track's chain at once instead of all replaying one.  Durations travel through a shared-memory
// tile [track][32 fragments]: filled with coalesced asynchronous copies one round ahead (a row per
// track, the rows shared out over the CTA's four warps), walked along the rows by warp 0 (lane =
// track), written back out along the columns by all four warps again.
constexpr int kAsmWarps = 8;
__global__ void __launch_bounds__(32 * kAsmWarps)
m3u8_assemble_tracks_kernel(const AssembleParams p)
{
    __shared__ __align__(16) double tile[2][32][33];      // [buffer][track][fragment], padded rows
    __shared__ int sn_tile[32][33];
    __shared__ __align__(16) longlong2 span_of[32];       // per track: {first fragment, fragments}
    __shared__ long long maxlen_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long n_tags_seen = (long long)p.out->n_tags;
    if ((long long)p.out->n_fragments > p.capacity || n_tags_seen > p.tag_capacity) return;
    // per-track state lives in warp 0: lane = track
    const long long t = (long long)blockIdx.x * 32 + lane;
    const long long none = 0x7fffffffffffffffll;
    long long off = 0, len = 0, e = 0, te = 0, next_tag = none, cur = 0;
    double total = 0.0;
    bool overflow = false;
    if (warp == 0) {
        const bool valid = t < p.n_tracks;
        const bool parsed = valid && (p.state == nullptr || p.state[t] == HLSMAP_TRACK_PARSED);
        off = valid ? p.track_off[t] : 0;
        len = parsed ? p.track_off[t + 1] - off : 0;
        if (valid) p.dur0[t] = len > 0 ? p.dur[off] : __longlong_as_double(0x7ff8000000000000ll);
        if (len > 0) {          // sequence tags of this lane's track (see m3u8_assemble_kernel)
            const long long tb = p.text_off[t];
            te = p.text_off[t + 1];
            long long z = n_tags_seen;
            if (t < n_tags_seen && p.tag_pos[t] >= tb && (t == 0 || p.tag_pos[t - 1] < tb)) {
                e = t;
            } else {
                while (e < z) {
                    const long long mid = (e + z) >> 1;
                    if (p.tag_pos[mid] < tb) e = mid + 1;
                    else z = mid;
                }
            }
            next_tag = (e < n_tags_seen && p.tag_pos[e] < te) ? p.tag_frag[e] : none;
        }
        long long m = len;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            const long long o = __shfl_xor_sync(0xffffffffu, m, d);
            m = o > m ? o : m;
        }
        span_of[lane] = make_longlong2(off, len);
        if (lane == 0) maxlen_s = m;
    }
    __syncthreads();
    const long long maxlen = maxlen_s;

    // fills buffer b with fragments [c, c + 32) of the 32 tracks: warp w takes rows 8w .. 8w+7
    auto fill = [&](int b, long long c) {
#pragma unroll
        for (int kk = 0; kk < 32 / kAsmWarps; kk++) {
            const int k = warp * (32 / kAsmWarps) + kk;
            const longlong2 sk = span_of[k];
            double *dst = &tile[b][k][lane];
            if (c + lane < sk.y) {
                const unsigned sa = (unsigned)__cvta_generic_to_shared(dst);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(p.dur + sk.x + c + lane) : "memory");
            } else {
                *dst = 0.0;
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    int buf = 0;
    if (maxlen > 0) fill(0, 0);
    for (long long c = 0; c < maxlen; c += 32) {
        if (c + 32 < maxlen) {
            fill(buf ^ 1, c + 32);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        if (warp == 0) {        // lane = track: walk the row in order
            const int n = len - c >= 32 ? 32 : (len > c ? (int)(len - c) : 0);
            double *row = tile[buf][lane];
            int *snrow = sn_tile[lane];
            if (next_tag > off + c + n - 1) {          // no tag inside: numbering runs straight through
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    const double d = row[j];           // 0.0 past the end: total + 0.0 == total
                    row[j] = total;
                    total += d;
                    snrow[j] = (int)(cur + j);
                }
                overflow |= n > 0 && cur + n - 1 > 0x7fffffffll;
                cur += n;
            } else {
                for (int j = 0; j < n; j++) {
                    while (next_tag <= off + c + j) {
                        cur = p.tag_val[e++];
                        next_tag = (e < n_tags_seen && p.tag_pos[e] < te) ? p.tag_frag[e] : none;
                    }
                    const double d = row[j];
                    row[j] = total;
                    total += d;
                    overflow |= cur > 0x7fffffffll;
                    snrow[j] = (int)cur;
                    cur++;
                }
            }
        }
        __syncthreads();
        // lane = fragment: write the columns out, a track at a time
#pragma unroll
        for (int kk = 0; kk < 32 / kAsmWarps; kk++) {
            const int k = warp * (32 / kAsmWarps) + kk;
            const longlong2 sk = span_of[k];
            if (c + lane < sk.y) {
                p.start[sk.x + c + lane] = tile[buf][k][lane];
                p.sn[sk.x + c + lane] = sn_tile[k][lane];
            }
        }
        __syncthreads();
        buf ^= 1;
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

cudaError_t launch_ingest_compact(const IngestParams &p, cudaStream_t s)
{
    if (p.n_spans < 1) return cudaSuccess;
    m3u8_block_scan_kernel<<<1, 1024, 0, s>>>(p);
    m3u8_span_base_kernel<<<p.n_blocks, 256, 0, s>>>(p);
    if (p.stage_dur != nullptr) m3u8_copy_kernel<<<(unsigned)((p.n_spans + 7) / 8), 256, 0, s>>>(p);
    m3u8_track_offsets_kernel<<<(unsigned)((p.n_tracks + 1 + 255) / 256), 256, 0, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_ingest_assemble(const AssembleParams &p, bool warp_per_track, cudaStream_t s)
{
    if (p.n_tracks <= 0) return cudaSuccess;
    static const bool old_warp = getenv("HLSMAP_ASSEMBLE_WARP") != nullptr;     // diagnostics: the warp-per-track form
    if (warp_per_track && old_warp) {
        const long long threads = (long long)p.n_tracks * 32;
        m3u8_assemble_kernel<32><<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(p);
    } else if (warp_per_track) {
        m3u8_assemble_tracks_kernel<<<(unsigned)((p.n_tracks + 31) / 32), 32 * kAsmWarps, 0, s>>>(p);
    } else if (old_warp) {
        m3u8_assemble_kernel<1><<<(unsigned)((p.n_tracks + 255) / 256), 256, 0, s>>>(p);
    } else {
        m3u8_assemble_short_kernel<<<(unsigned)((p.n_tracks + 32 * kShortWarps - 1) / (32 * kShortWarps)),
                                     32 * kShortWarps, 0, s>>>(p);
    }
    return cudaGetLastError();
}

}  // namespace hlsmap

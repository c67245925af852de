// decimal.cuh -- exact decimal text -> binary64, as JavaScript's parseFloat rounds it.
//
// The playlist parser of hls.js reads `#EXTINF:` durations with parseFloat on a run of
// [0-9.] characters; a fragment table built on the device has to hold the
// very same doubles, so the conversion below is correctly rounded (round to nearest, ties to
// even) wherever it answers, and says "unsupported" instead of guessing where it cannot:
//   * at most 19 integer digits and a decimal exponent no smaller than -38 after the first 19
//     significant digits;
//   * with more than 19 significant digits, only when the dropped tail cannot change the
//     rounding.
// Plain integer arithmetic: compiles for the host too (tests/test_decimal_host.py).
#pragma once
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define HLSMAP_HD __host__ __device__
#else
#define HLSMAP_HD
#endif

namespace hlsmap {

struct DecimalScan {
    unsigned long long w;  // the first (up to 19) significant digits
    int exp10;             // value = w * 10^exp10, plus whatever `sticky` stands for
    int run;               // characters of the [0-9.]+ run (the regex capture)
    bool sticky;           // a non-zero digit beyond the 19th was dropped
    bool is_number;        // the run starts with a StrDecimalLiteral (else parseFloat gives NaN)
};

// Reads the run of [0-9.] starting at p (never past lim) and the StrDecimalLiteral prefix of
// it: digits, at most one '.', digits.  `get(pos)` returns the byte at an absolute position.
template <class Get>
HLSMAP_HD inline DecimalScan scan_decimal(const Get &get, long long p, long long lim)
{
    DecimalScan r;
    r.w = 0;
    r.exp10 = 0;
    r.run = 0;
    r.sticky = false;
    r.is_number = false;
    int ndig = 0;          // significant digits taken into w
    bool seen_dot = false, in_prefix = true;
    for (long long q = p; q < lim; q++) {
        const int c = get(q);
        const bool digit = c >= '0' && c <= '9';
        if (!digit && c != '.') break;
        r.run++;
        if (!in_prefix) continue;
        if (c == '.') {
            if (seen_dot) in_prefix = false;  // a second '.' ends the literal, not the run
            else seen_dot = true;
            continue;
        }
        r.is_number = true;
        const int d = c - '0';
        if (ndig == 0 && d == 0) {            // leading zero: only moves the exponent
            if (seen_dot) r.exp10--;
            continue;
        }
        if (ndig < 19) {
            r.w = r.w * 10ull + (unsigned long long)d;
            ndig++;
            if (seen_dot) r.exp10--;
        } else {
            if (d) r.sticky = true;
            if (!seen_dot) r.exp10++;
        }
    }
    if (r.w == 0) r.exp10 = 0;
    return r;
}

#if defined(__CUDACC__) || defined(__SIZEOF_INT128__)
HLSMAP_HD inline int clz64(unsigned long long x)
{
#ifdef __CUDA_ARCH__
    return __clzll((long long)x);
#else
    return __builtin_clzll(x);
#endif
}

// The 64 leading quotient bits of w / 10^k for k <= 19 (the divisor fits 64 bits), the binary
// exponent of the quotient and whether a remainder is left.  One floating-point division gives
// the quotient to within 2^13 units of its last bit; the difference is then settled exactly in
// integers (a 128-bit remainder, one more small estimate, at most a step or two of fix-up) --
// a few dozen instructions where the subtract-and-shift division below takes over a thousand.
// Returns false if an estimate is ever further off than the analysis allows (the caller then
// divides the slow way; not observed).
HLSMAP_HD inline bool div_pow10_estimate(unsigned long long w, int k, unsigned long long *q_out, int *e_out,
                                         bool *rem_out)
{
    unsigned long long d = 1;
    for (int i = 0; i < k; i++) d *= 10ull;
    const int la = 63 - clz64(w), ld = 63 - clz64(d);
    const int e0 = la - ld - (((w << (63 - la)) >= (d << (63 - ld))) ? 0 : 1);    // floor(log2(w / d))
    const int t = 63 - e0;                                                        // w * 2^t / d in [2^63, 2^64)
    const unsigned __int128 num = (unsigned __int128)w << t;
#ifdef __CUDA_ARCH__
    const double x0 = __ddiv_rn(__ull2double_rn(w), __ull2double_rn(d));
#else
    const double x0 = (double)w / (double)d;
#endif
    unsigned long long xb;
    memcpy(&xb, &x0, sizeof xb);
    const int sh = (int)((xb >> 52) & 0x7ffull) - 1075 + t;                       // x0 * 2^t = m * 2^sh
    const unsigned long long m = (xb & ((1ull << 52) - 1ull)) | (1ull << 52);
    if (sh < 8 || sh > 12) return false;
    unsigned long long q = sh == 12 ? ~0ull << 11 : m << sh;                      // (m << 12 would be 2^64)
    __int128 r = (__int128)(num - (unsigned __int128)q * d);                      // |r| < 2^14 * d
    // r / d, estimated: r itself fits 64 bits next to a small divisor, r / 2^13 does next to a large one
    const bool wide = d >= (1ull << 48);
    const long long rs = wide ? (long long)(r >> 13) : (long long)r;
    const double scale = wide ? 8192.0 : 1.0;
#ifdef __CUDA_ARCH__
    const long long delta = __double2ll_rd(__ddiv_rn(__dmul_rn(__ll2double_rn(rs), scale), __ull2double_rn(d)));
#else
    const double de = (double)rs * scale / (double)d;
    long long delta = (long long)de;
    if ((double)delta > de) delta--;
#endif
    if (delta > (1ll << 20) || delta < -(1ll << 20)) return false;
    r -= (__int128)delta * (__int128)(unsigned __int128)d;
    q += (unsigned long long)delta;
    for (int i = 0; i < 8 && r < 0; i++) {
        r += (__int128)(unsigned __int128)d;
        q--;
    }
    for (int i = 0; i < 8 && r >= (__int128)(unsigned __int128)d; i++) {
        r -= (__int128)(unsigned __int128)d;
        q++;
    }
    if (r < 0 || r >= (__int128)(unsigned __int128)d || !(q >> 63)) return false;
    *q_out = q;
    *e_out = e0;
    *rem_out = r != 0;
    return true;
}

// 64 quotient bits (the first of them set), the quotient's binary exponent and a sticky flag ->
// binary64, round to nearest, ties to even.
HLSMAP_HD inline double round_quotient(unsigned long long q, int e, bool sticky)
{
    unsigned long long m = q >> 11;
    const unsigned long long rest = q & 0x7ffull;
    if (rest > 0x400ull || (rest == 0x400ull && (sticky || (m & 1ull)))) m++;
    if (m == (1ull << 53)) {
        m >>= 1;
        e++;
    }
    const unsigned long long bits = ((unsigned long long)(e + 1023) << 52) | (m & ((1ull << 52) - 1ull));
    double out;
    memcpy(&out, &bits, sizeof out);
    return out;
}

// w / 10^k (0 <= k <= 38, w != 0) rounded to binary64; `above` = the true value lies
// strictly above w / 10^k by less than the quotient's last kept bit can resolve.  ESTIMATE = false
// leaves the estimated division out (callers for which this path is rare and registers are dear).
template <bool ESTIMATE = true>
HLSMAP_HD inline double div_pow10_rounded(unsigned long long w, int k, bool above)
{
    if (ESTIMATE && k <= 19) {
        unsigned long long qf;
        int ef;
        bool rem;
        if (div_pow10_estimate(w, k, &qf, &ef, &rem)) return round_quotient(qf, ef, above || rem);
    }
    unsigned __int128 d = 1, a = w;
    for (int i = 0; i < k; i++) d *= 10u;
    int e = 0;  // value = (a / d) * 2^e with d <= a < 2d
    if (a >= d) {
        while ((d << 1) <= a) {
            d <<= 1;
            e++;
        }
    } else {
        while (a < d) {
            a <<= 1;
            e--;
        }
    }
    unsigned long long q = 0;  // 64 quotient bits, the first of them set
    for (int i = 0; i < 64; i++) {
        q <<= 1;
        if (a >= d) {
            a -= d;
            q |= 1ull;
        }
        a <<= 1;
    }
    return round_quotient(q, e, above || a != 0);
}

// false = outside what this converter answers exactly (the caller reports it, loudly).
template <bool ESTIMATE = true>
HLSMAP_HD inline bool decimal_to_double(const DecimalScan &s, double *out)
{
    if (s.w == 0) {
        *out = 0.0;
        return true;
    }
    if (s.exp10 > 0 || s.exp10 < -38) return false;
    const int k = -s.exp10;
    if (!s.sticky && s.w <= (1ull << 53) && k <= 22) {
        // both operands are exact doubles (10^k = 5^k * 2^k, 5^22 < 2^53): one IEEE division
        unsigned long long p5 = 1;
        for (int i = 0; i < k; i++) p5 *= 5ull;
        double den = (double)p5;
        for (int i = 0; i < k; i++) den *= 2.0;
        *out = (double)s.w / den;
        return true;
    }
    const double lo = div_pow10_rounded<ESTIMATE>(s.w, k, s.sticky);
    if (!s.sticky) {
        *out = lo;
        return true;
    }
    const double hi = div_pow10_rounded<ESTIMATE>(s.w + 1ull, k, false);  // w < 10^19: no overflow
    if (lo != hi) return false;
    *out = lo;
    return true;
}
#endif

// ---- the same scan on 16 bytes held in two registers (SWAR) -------------------------------------
// x0 holds text bytes 0..7, x1 bytes 8..15, first byte lowest; bytes past the end of the text
// must be zero.  Answers what scan_decimal() answers when the [0-9.]+ run ends inside the 16
// bytes and neither side of the '.' has more than 8 digits; returns false otherwise (the caller
// then walks the bytes).  run = characters of the run; value = w * 10^-nf.
HLSMAP_HD inline unsigned long long swar_nondigit(unsigned long long x)
{
    const unsigned long long t = x ^ 0x3030303030303030ull;
    return (((t & 0x7f7f7f7f7f7f7f7full) + 0x7676767676767676ull) | t) & 0x8080808080808080ull;
}
HLSMAP_HD inline unsigned long long swar_is_byte(unsigned long long x, unsigned char c)
{
    const unsigned long long y = x ^ (0x0101010101010101ull * c);
    return ((((y & 0x7f7f7f7f7f7f7f7full) + 0x7f7f7f7f7f7f7f7full) | y) & 0x8080808080808080ull) ^ 0x8080808080808080ull;
}
HLSMAP_HD inline int swar_first(unsigned long long m0, unsigned long long m1)   // first flagged byte, 16 if none
{
#ifdef __CUDA_ARCH__
    if (m0) return (__ffsll((long long)m0) - 1) >> 3;
    if (m1) return 8 + ((__ffsll((long long)m1) - 1) >> 3);
#else
    if (m0) return __builtin_ctzll(m0) >> 3;
    if (m1) return 8 + (__builtin_ctzll(m1) >> 3);
#endif
    return 16;
}
HLSMAP_HD inline unsigned swar_parse8(unsigned long long v)   // eight ASCII digits (or zero bytes), first lowest
{
    v = ((v & 0x0f0f0f0f0f0f0f0full) * 2561ull) >> 8;
    v = ((v & 0x00ff00ff00ff00ffull) * 6553601ull) >> 16;
    return (unsigned)(((v & 0x0000ffff0000ffffull) * 42949672960001ull) >> 32);
}
HLSMAP_HD inline unsigned long long swar_bytes_from(unsigned long long x0, unsigned long long x1, int at)
{   // the eight bytes starting at byte `at` (0..16) of the 16
    if (at == 0) return x0;
    if (at < 8) return (x0 >> (8 * at)) | (x1 << (64 - 8 * at));
    if (at == 8) return x1;
    if (at < 16) return x1 >> (8 * (at - 8));
    return 0ull;
}
HLSMAP_HD inline bool swar_scan_decimal16(unsigned long long x0, unsigned long long x1, int *run, unsigned long long *w,
                                          int *nf, bool *is_number)
{
    const unsigned long long d0 = swar_is_byte(x0, '.'), d1 = swar_is_byte(x1, '.');
    const int pe = swar_first(swar_nondigit(x0) & ~d0, swar_nondigit(x1) & ~d1);
    if (pe == 16) return false;                        // the run goes on past the window
    *run = pe;
    int pd = swar_first(d0, d1);                       // first '.', if inside the run
    int ni = pe, nfrac = 0;
    if (pd < pe) {
        const unsigned long long e0 = d0 & (d0 - 1), e1 = d0 ? d1 : (d1 & (d1 - 1));   // dots after the first
        int le = swar_first(e0, e1);
        if (le > pe) le = pe;
        ni = pd;
        nfrac = le - pd - 1;
    }
    *is_number = ni + nfrac > 0;
    if (ni > 8 || nfrac > 8) return false;
    const unsigned ip = ni ? swar_parse8(x0 << (8 * (8 - ni))) : 0u;
    unsigned fp = 0;
    if (nfrac) fp = swar_parse8(swar_bytes_from(x0, x1, pd + 1) << (8 * (8 - nfrac)));
    unsigned p10 = 1u;                                  // 10^nfrac, nfrac <= 8, without a loop
    if (nfrac & 1) p10 *= 10u;
    if (nfrac & 2) p10 *= 100u;
    if (nfrac & 4) p10 *= 10000u;
    if (nfrac & 8) p10 *= 100000000u;
    *w = (unsigned long long)ip * p10 + fp;
    *nf = nfrac;
    return true;
}

// ---- JSON numbers: -? int frac? exp?  (RFC 8259 / ECMA-404, what JSON.parse accepts) ---------
struct JsonNumber {
    unsigned long long w;  // first (up to 19) significant digits
    int exp10;
    int length;            // characters consumed; 0 = not a JSON number
    bool sticky, negative;
};

template <class Get>
HLSMAP_HD inline JsonNumber scan_json_number(const Get &get, long long p, long long lim)
{
    JsonNumber r;
    r.w = 0;
    r.exp10 = 0;
    r.length = 0;
    r.sticky = false;
    r.negative = false;
    long long q = p;
    if (q < lim && get(q) == '-') {
        r.negative = true;
        q++;
    }
    int ndig = 0, int_digits = 0;
    const long long int_start = q;
    while (q < lim) {                       // int: "0" or [1-9][0-9]*
        const int c = get(q);
        if (c < '0' || c > '9') break;
        if (int_digits == 1 && get(int_start) == '0') return r;   // leading zero
        int_digits++;
        const int d = c - '0';
        if (ndig > 0 || d > 0) {
            if (ndig < 19) {
                r.w = r.w * 10ull + (unsigned long long)d;
                ndig++;
            } else {
                if (d) r.sticky = true;
                r.exp10++;
            }
        }
        q++;
    }
    if (int_digits == 0) return r;
    if (q < lim && get(q) == '.') {
        q++;
        int frac_digits = 0;
        while (q < lim) {
            const int c = get(q);
            if (c < '0' || c > '9') break;
            frac_digits++;
            const int d = c - '0';
            if (ndig > 0 || d > 0) {
                if (ndig < 19) {
                    r.w = r.w * 10ull + (unsigned long long)d;
                    ndig++;
                    r.exp10--;
                } else if (d) {
                    r.sticky = true;
                }
            } else {
                r.exp10--;
            }
            q++;
        }
        if (frac_digits == 0) return r;
    }
    if (q < lim && (get(q) == 'e' || get(q) == 'E')) {
        q++;
        bool eneg = false;
        if (q < lim && (get(q) == '+' || get(q) == '-')) {
            eneg = get(q) == '-';
            q++;
        }
        int e = 0, edigits = 0;
        while (q < lim) {
            const int c = get(q);
            if (c < '0' || c > '9') break;
            edigits++;
            if (e < 100000) e = e * 10 + (c - '0');
            q++;
        }
        if (edigits == 0) return r;
        r.exp10 += eneg ? -e : e;
    }
    if (r.w == 0) r.exp10 = 0;
    r.length = (int)(q - p);
    return r;
}

#if defined(__CUDACC__) || defined(__SIZEOF_INT128__)
// |value| = w * 10^exp10 (+ sticky) -> binary64, or false when outside what is answered exactly:
// here exp10 may also be positive, up to 19 (w * 10^19 < 2^128).
HLSMAP_HD inline bool json_number_to_double(const JsonNumber &n, double *out)
{
    double mag = 0.0;
    if (n.w == 0 || n.exp10 <= 0) {
        DecimalScan s;
        s.w = n.w;
        s.exp10 = n.exp10;
        s.run = 1;
        s.sticky = n.sticky;
        s.is_number = true;
        if (!decimal_to_double(s, &mag)) return false;
    } else {
        if (n.sticky || n.exp10 > 19) return false;
        unsigned __int128 v = n.w;
        for (int i = 0; i < n.exp10; i++) v *= 10u;
        int top = 127;
        while (!((v >> top) & 1)) top--;
        if (top <= 52) {
            mag = (double)(unsigned long long)v;
        } else {
            const int sh = top - 52;
            unsigned long long m = (unsigned long long)(v >> sh);
            const unsigned __int128 rest = v & ((((unsigned __int128)1) << sh) - 1);
            const unsigned __int128 half = ((unsigned __int128)1) << (sh - 1);
            int e = top;
            if (rest > half || (rest == half && (m & 1ull))) m++;
            if (m == (1ull << 53)) {
                m >>= 1;
                e++;
            }
            const unsigned long long bits = ((unsigned long long)(e + 1023) << 52) | (m & ((1ull << 52) - 1ull));
            memcpy(&mag, &bits, sizeof mag);
        }
    }
    *out = n.negative ? -mag : mag;
    return true;
}
#endif

}  // namespace hlsmap

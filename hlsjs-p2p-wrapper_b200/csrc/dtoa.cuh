// dtoa.cuh -- binary64 -> the shortest decimal that reads back as the same double, and its
// rendering as ECMAScript's Number::toString (what JSON.stringify prints for `time`).
//
// JSON.stringify(segmentView) (test/segment-view.js:5-11) serialises `time`
// with Number::toString: the fewest digits k such that the decimal still denotes the same Number,
// the closest such decimal, then one of four layouts depending on the exponent.  The digit
// generation below is the Schubfach scheme (R. Giulietti, "The Schubfach way to render doubles"):
// one table look-up of a 126-bit approximation g of a power of ten, three 64x126-bit products
// rounded to odd, and a handful of comparisons; the table is generated exactly by
// scripts/gen_pow10_table.py.  Integer arithmetic only; compiles for the host as well, where
// tests/test_json_host.py checks it against glibc's printf/strtod search and CPython's repr.
#pragma once
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define HLSMAP_HD __host__ __device__
#else
#define HLSMAP_HD
#endif

namespace hlsmap {

constexpr int kPow10TableMin = -324;   // first k of the table
constexpr int kPow10TableLen = 617;

struct Shortest {
    unsigned long long digits;  // decimal significand without trailing zeros (0 for +-0)
    int exp10;                  // value = digits * 10^exp10
    int ndigits;                // decimal digits of `digits` (1 for 0)
};

HLSMAP_HD inline unsigned long long mulhi64(unsigned long long a, unsigned long long b)
{
#ifdef __CUDA_ARCH__
    return __umul64hi(a, b);
#else
    return (unsigned long long)(((unsigned __int128)a * b) >> 64);
#endif
}

// g * cp / 2^127 rounded to odd (g = g1 * 2^63 + g0).
HLSMAP_HD inline unsigned long long round_to_odd(unsigned long long g1, unsigned long long g0,
                                                 unsigned long long cp)
{
    const unsigned long long x1 = mulhi64(g0, cp);
    const unsigned long long y0 = g1 * cp;
    const unsigned long long y1 = mulhi64(g1, cp);
    const unsigned long long z = (y0 >> 1) + x1;
    const unsigned long long w = y1 + (z >> 63);
    const unsigned long long m63 = 0x7fffffffffffffffull;
    return w | (((z & m63) + m63) >> 63);
}

HLSMAP_HD inline int decimal_length(unsigned long long v)
{
    int n = 1;
    unsigned long long p = 10;
    while (n < 20 && v >= p) {
        n++;
        if (n < 20) p *= 10;
    }
    return n;
}

// Shortest digits of a finite, non-zero |x| given as raw bits (sign ignored).  `table` is the
// {g1, g0} array of pow10_table.inc.
HLSMAP_HD inline Shortest shortest_digits(unsigned long long bits, const unsigned long long (*table)[2])
{
    const unsigned long long t = bits & ((1ull << 52) - 1ull);
    const int bq = (int)((bits >> 52) & 0x7ffu);
    unsigned long long c;
    int q;
    Shortest r;
    if (bq > 0) {
        c = (1ull << 52) | t;
        q = bq - 1075;
        if (q <= 0 && q > -53) {                    // an integer below 2^53 prints as itself
            const int mq = -q;
            const unsigned long long f = c >> mq;
            if ((f << mq) == c) {
                unsigned long long d = f;
                int e = 0;
                while (d % 10ull == 0ull) {
                    d /= 10ull;
                    e++;
                }
                r.digits = d;
                r.exp10 = e;
                r.ndigits = decimal_length(d);
                return r;
            }
        }
    } else {                                        // subnormal
        q = -1074;
        c = t;
    }
    const int out = (int)(c & 1ull);
    const unsigned long long cb = c << 2, cbr = cb + 2ull;
    unsigned long long cbl;
    int k;
    if (c != (1ull << 52) || q == -1074) {          // neighbours equally spaced
        cbl = cb - 2ull;
        k = (int)(((long long)q * 661971961083ll) >> 41);
    } else {                                        // a power of two: the lower neighbour is closer
        cbl = cb - 1ull;
        k = (int)(((long long)q * 661971961083ll - 274743187321ll) >> 41);
    }
    const int h = q + (int)(((long long)(-k) * 913124641741ll) >> 38) + 2;
    const unsigned long long g1 = table[k - kPow10TableMin][0], g0 = table[k - kPow10TableMin][1];
    const unsigned long long vb = round_to_odd(g1, g0, cb << h);
    const unsigned long long vbl = round_to_odd(g1, g0, cbl << h);
    const unsigned long long vbr = round_to_odd(g1, g0, cbr << h);
    const unsigned long long s = vb >> 2;
    unsigned long long f;
    int e = k;
    bool chosen = false;
    {   // a decimal one digit shorter, if one fits (the interval is narrower than 10^(k+1), so at
        // most one multiple of it lies inside; taken whatever the size of s, which is what makes
        // tiny subnormals print with a single digit as ECMAScript wants)
        const unsigned long long sp10 = 10ull * (s / 10ull), tp10 = sp10 + 10ull;
        const bool upin = vbl + (unsigned long long)out <= (sp10 << 2);
        const bool wpin = (tp10 << 2) + (unsigned long long)out <= vbr;
        if (upin != wpin) {
            f = upin ? sp10 : tp10;
            chosen = true;
        }
    }
    if (!chosen) {
        const unsigned long long tt = s + 1ull;
        const bool uin = vbl + (unsigned long long)out <= (s << 2);
        const bool win = (tt << 2) + (unsigned long long)out <= vbr;
        if (uin != win) {
            f = uin ? s : tt;
        } else {                                    // both or neither: the closer one, ties to even
            const long long cmp = (long long)(vb - ((s + tt) << 1));
            f = (cmp < 0 || (cmp == 0 && (s & 1ull) == 0ull)) ? s : tt;
        }
    }
    while (f % 10ull == 0ull) {
        f /= 10ull;
        e++;
    }
    r.digits = f;
    r.exp10 = e;
    r.ndigits = decimal_length(f);
    return r;
}

// Characters Number::toString needs for these digits (without a sign).
HLSMAP_HD inline int js_number_length(const Shortest &s)
{
    const int k = s.ndigits, n = s.exp10 + k;       // value = 0.d1..dk * 10^n
    if (k <= n && n <= 21) return n;
    if (0 < n && n <= 21) return k + 1;
    if (-6 < n && n <= 0) return 2 - n + k;
    int ee = n - 1 < 0 ? 1 - n : n - 1, el = 1;
    while (ee >= 10) {
        ee /= 10;
        el++;
    }
    return (k == 1 ? 1 : k + 1) + 2 + el;           // digits [.digits] e sign exponent
}

// Writes Number::toString of the digits at dst (no sign, no terminator); returns the length.
// Digits are peeled off the low end and stored straight at their final position (no staging
// array), in two 32-bit halves so that the divisions are 32-bit ones.
HLSMAP_HD inline int js_number_write(const Shortest &s, char *dst)
{
    const int k = s.ndigits, n = s.exp10 + k;
    // layout: where digit i (0 = most significant) goes, and what surrounds the digits
    int shift_from = k, lead = 0, len;       // digit i lands at lead + i + (i >= shift_from)
    if (k <= n && n <= 21) {
        for (int i = k; i < n; i++) dst[i] = '0';
        len = n;
    } else if (0 < n && n <= 21) {
        shift_from = n;
        dst[n] = '.';
        len = k + 1;
    } else if (-6 < n && n <= 0) {
        dst[0] = '0';
        dst[1] = '.';
        for (int i = 0; i < -n; i++) dst[2 + i] = '0';
        lead = 2 - n;
        len = lead + k;
    } else {
        int at = k;
        if (k > 1) {
            shift_from = 1;
            dst[1] = '.';
            at = k + 1;
        }
        dst[at++] = 'e';
        int ee = n - 1;
        dst[at++] = ee < 0 ? '-' : '+';
        if (ee < 0) ee = -ee;
        if (ee >= 100) dst[at++] = (char)('0' + ee / 100);
        if (ee >= 10) dst[at++] = (char)('0' + (ee / 10) % 10);
        dst[at++] = (char)('0' + ee % 10);
        len = at;
    }
    unsigned lo = (unsigned)(s.digits % 1000000000ull);
    unsigned long long hi = s.digits / 1000000000ull;      // < 10^11: at most two more 9-digit groups
    int i = k - 1;
    for (int g = 0; g < 3 && i >= 0; g++) {
        for (int j = 0; j < 9 && i >= 0; j++, i--) {
            dst[lead + i + (i >= shift_from ? 1 : 0)] = (char)('0' + lo % 10u);
            lo /= 10u;
        }
        lo = (unsigned)(hi % 1000000000ull);
        hi /= 1000000000ull;
    }
    return len;
}

}  // namespace hlsmap

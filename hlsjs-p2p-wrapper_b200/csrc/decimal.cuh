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
// w / 10^k (0 <= k <= 38, w != 0) rounded to binary64; `above` = the true value lies
// strictly above w / 10^k by less than the quotient's last kept bit can resolve.
HLSMAP_HD inline double div_pow10_rounded(unsigned long long w, int k, bool above)
{
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
    const bool sticky = above || a != 0;
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

// false = outside what this converter answers exactly (the caller reports it, loudly).
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
    const double lo = div_pow10_rounded(s.w, k, s.sticky);
    if (!s.sticky) {
        *out = lo;
        return true;
    }
    const double hi = div_pow10_rounded(s.w + 1ull, k, false);  // w < 10^19: no overflow
    if (lo != hi) return false;
    *out = lo;
    return true;
}
#endif

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

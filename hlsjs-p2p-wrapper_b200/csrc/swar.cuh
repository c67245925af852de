// swar.cuh -- eight text bytes at a time (SIMD within a register): digit runs and their value.
//
// A little-endian 64-bit word holds text byte k in bits [8k, 8k + 8).  Used by the JSON reader to
// take an int32 column value in two or three word operations instead of a byte loop.  Plain integer
// arithmetic: compiles for the host too (tests/test_json_host.py checks it there exhaustively enough).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define HLSMAP_HD __host__ __device__
#else
#define HLSMAP_HD
#endif

namespace hlsmap {

// How many of the word's leading bytes (from byte 0 on) are ASCII digits: 0 .. 8.
HLSMAP_HD inline int leading_digits8(unsigned long long v)
{
    const unsigned long long t = v ^ 0x3030303030303030ull;          // a digit byte becomes 0 .. 9
    // high bit of a byte set <=> that byte is > 9 (no carry crosses a byte: the addend meets 7-bit values)
    const unsigned long long bad = (((t & 0x7f7f7f7f7f7f7f7full) + 0x7676767676767676ull) | t) & 0x8080808080808080ull;
    if (bad == 0) return 8;
#ifdef __CUDA_ARCH__
    return (__ffsll((long long)bad) - 1) >> 3;
#else
    return __builtin_ctzll(bad) >> 3;
#endif
}

// Value of the first nd (1 .. 8) bytes of the word, all of them ASCII digits, first byte most significant.
HLSMAP_HD inline unsigned digits8_value(unsigned long long v, int nd)
{
    unsigned long long t = (v ^ 0x3030303030303030ull) << (8 * (8 - nd));   // digits in the top nd bytes, zeros below
    t = t * 10 + (t >> 8);                                                  // pairs:  d0 d1 -> 10 d0 + d1 (odd bytes)
    const unsigned long long mask = 0x000000ff000000ffull;
    t = ((t & mask) * (100ull + (1000000ull << 32)) + ((t >> 16) & mask) * (1ull + (10000ull << 32))) >> 32;
    return (unsigned)t;
}

}  // namespace hlsmap

#include "decimal.cuh"

namespace hlsmap {

// An int32 exactly as JSON.stringify prints one (an optional '-', 1 .. 10 digits, no leading zero,
// not followed by '.', 'e', 'E' or another digit), read from `get8(pos)` = the 8 text bytes at pos.
// Nothing at or beyond `lim` counts.  false: not that shape -- the general reader decides.
template <class Get8>
HLSMAP_HD inline bool swar_int32(const Get8 &get8, long long &q, long long lim, int *out)
{
    unsigned long long v = get8(q);
    const bool neg = q < lim && (v & 0xffull) == '-';
    const long long r = q + (neg ? 1 : 0);
    if (neg) v = get8(r);
    const long long avail = lim - r;
    int nd = leading_digits8(v);
    if (nd > avail) nd = (int)avail;
    if (nd <= 0) return false;
    unsigned long long val = digits8_value(v, nd);
    int total = nd;
    if (nd == 8) {
        const unsigned long long v2 = get8(r + 8);
        int nd2 = leading_digits8(v2);
        if (nd2 > avail - 8) nd2 = (int)(avail - 8);
        if (nd2 > 2) return false;
        if (nd2 > 0) {
            val = val * (nd2 == 1 ? 10ull : 100ull) + digits8_value(v2, nd2);
            total += nd2;
        }
    }
    const int nx = r + total < lim ? (int)(get8(r + total) & 0xffull) : 0;
    if (nx == '.' || nx == 'e' || nx == 'E') return false;
    if (total > 1 && (v & 0xffull) == '0') return false;
    if (val > (neg ? 2147483648ull : 2147483647ull)) return false;
    *out = (int)(neg ? 0ll - (long long)val : (long long)val);
    q = r + total;
    return true;
}

// digits [. digits] without an exponent: at most 7 integer digits, at most 15 fraction digits, at
// most 19 digits in all.  Correctly rounded: up to 15 significant digits one IEEE division of two
// exact doubles, beyond that the exact integer division of decimal.cuh.  pow10[k] = 10^k for
// k = 0 .. 22 as doubles (all exact), ipow10[k] = 10^k for k = 0 .. 19 as integers: the caller's
// tables (constant memory on the device).  false: the caller's other readers decide.
template <class Get8>
HLSMAP_HD inline bool swar_plain_decimal(const Get8 &get8, long long &q, long long lim, double *out,
                                         const double *pow10, const unsigned long long *ipow10)
{
    unsigned long long v = get8(q);
    const bool neg = q < lim && (v & 0xffull) == '-';
    long long r = q + (neg ? 1 : 0);
    if (neg) v = get8(r);
    int ni = leading_digits8(v);
    if (ni > lim - r) ni = (int)(lim - r);
    if (ni <= 0 || ni >= 8 || (ni > 1 && (v & 0xffull) == '0')) return false;
    unsigned long long w = digits8_value(v, ni);
    r += ni;
    int frac = 0;
    if (r < lim && ((v >> (8 * ni)) & 0xffull) == '.') {
        r++;
        const unsigned long long f1 = get8(r);
        int n1 = leading_digits8(f1);
        if (n1 > lim - r) n1 = (int)(lim - r);
        if (n1 <= 0) return false;
        unsigned long long fw = digits8_value(f1, n1);
        frac = n1;
        r += n1;
        if (n1 == 8) {
            const unsigned long long f2 = get8(r);
            int n2 = leading_digits8(f2);
            if (n2 > lim - r) n2 = (int)(lim - r);
            if (n2 >= 8) return false;
            if (n2 > 0) {
                fw = fw * ipow10[n2] + digits8_value(f2, n2);
                frac += n2;
                r += n2;
            }
        }
        if (ni + frac > 19) return false;           // the product below would leave 64 bits
        w = w * ipow10[frac] + fw;
    }
    const int nx = r < lim ? (int)(get8(r) & 0xffull) : 0;
    if (nx == 'e' || nx == 'E' || nx == '.' || (nx >= '0' && nx <= '9')) return false;
    double mag;
    if (w == 0) mag = 0.0;
    else if (w < 1000000000000000ull) {             // at most 15 significant digits
#ifdef __CUDA_ARCH__
        mag = frac ? __ddiv_rn((double)w, pow10[frac]) : (double)w;
#else
        mag = frac ? (double)w / pow10[frac] : (double)w;
#endif
    } else if (w < 10000000000000000000ull) {
        mag = div_pow10_rounded(w, frac, false);
    } else {
        return false;
    }
    *out = neg ? -mag : mag;
    q = r;
    return true;
}

}  // namespace hlsmap

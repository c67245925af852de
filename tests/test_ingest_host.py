"""CPU-side checks of the playlist ingest: the oracle's restatement of hls.js
0.5.46's parseLevelPlaylist on hand-computed cases, and the product's integer-only decimal
converter (csrc/decimal.cuh, compiled here for the host) against glibc strtod and Python float()."""
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle
from oracle import pyref
from hlsjs_p2p_wrapper_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "hlsjs-p2p-wrapper_b200", "csrc")


def _parse(text):
    return oracle.parse_m3u8(text, [0, len(text)])


def test_oracle_known_answers():
    r = _parse(b"#EXTM3U\n#EXT-X-TARGETDURATION:10\n#EXT-X-MEDIA-SEQUENCE:2680\n#EXTINF:9.9,\nseg2680.ts\n"
               b"#EXTINF:10,\nseg2681.ts\n#EXTINF:0.1,\nseg2682.ts\n")
    assert list(r["sn"]) == [2680, 2681, 2682] and list(r["live"]) == [1]
    assert list(r["duration"]) == [9.9, 10.0, 0.1] and list(r["start"]) == [0.0, 9.9, 9.9 + 10.0]
    # the corners, spelled out (see synth._NASTY): hidden tags, parseFloat prefixes, NaN runs
    r = _parse(synth._NASTY[0])
    assert list(r["duration"]) == [10.0, 0.5, 1.2] and list(r["sn"]) == [7, 8, 9] and list(r["live"]) == [0]
    r = _parse(synth._NASTY[1])
    assert list(r["duration"]) == [3.0, 2.0, 4.0, 5.0] and list(r["sn"]) == [0, 1, 12, 13] and list(r["live"]) == [0]
    r = _parse(synth._NASTY[2])
    assert list(r["duration"]) == [6.0, 0.0, 8.0, 9.98458] and list(r["live"]) == [1]
    r = _parse(synth._NASTY[3])
    assert list(r["duration"]) == [10.5, float("0.1234567890123456789"), 123456789012345678.0,
                                   0.30000000000000004, 9007199254740992.0]
    assert list(r["sn"]) == list(range(2147483640, 2147483645)) and not r["bad"]
    r = _parse(synth._NASTY[6])
    assert list(r["sn"]) == [5, 6, 100, 101]
    assert _parse(b"#EXT-X-MEDIA-SEQUENCE:2147483647\n#EXTINF:1,\n#EXTINF:1,\n")["bad"]


def test_oracle_start_is_the_sequential_sum():
    rng = np.random.default_rng(1)
    d = rng.integers(1, 10 ** 6, 500) / 1000.0
    text = "".join("#EXTINF:%r,\nx\n" % float(x) for x in d).encode()
    r = _parse(text)
    acc, exp = 0.0, []
    for x in d:
        exp.append(acc)
        acc += float(x)
    assert np.array_equal(r["duration"], d) and list(r["start"]) == exp


def test_oracle_threads_agree():
    base, state, texts = synth.m3u8_corpus(n_playlists=150, unparsed_every=9)
    off = np.concatenate([[0], np.cumsum([len(t) for t in texts])])
    a = oracle.parse_m3u8(b"".join(texts), off, state, threads=1)
    b = oracle.parse_m3u8(b"".join(texts), off, state, threads=5)
    for k in ("track_off", "live", "sn", "start", "duration"):
        assert np.array_equal(a[k], b[k])
    assert not np.diff(a["track_off"])[state != 2].any()


HARNESS = r"""
#include "decimal.cuh"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
using namespace hlsmap;
struct G { const char *s; int operator()(long long p) const { return (unsigned char)s[p]; } };
int main() {
    char line[256];
    while (fgets(line, sizeof line, stdin)) {
        size_t n = strlen(line);
        while (n && (line[n - 1] == '\n' || line[n - 1] == '\r')) line[--n] = 0;
        G g{line};
        DecimalScan d = scan_decimal(g, 0, (long long)n);
        double x = 0;
        if (!d.is_number) { printf("nan %d\n", d.run); continue; }
        if (!decimal_to_double(d, &x)) { printf("unsupported %d\n", d.run); continue; }
        unsigned long long b; memcpy(&b, &x, 8);
        char *end; double e = strtod(line, &end);
        unsigned long long eb; memcpy(&eb, &e, 8);
        printf("%016llx %016llx %d\n", b, eb, d.run);
    }
}
"""


def test_decimal_converter_matches_strtod_and_python(tmp_path):
    src, exe = tmp_path / "h.cpp", tmp_path / "h"
    src.write_text(HARNESS)
    subprocess.check_call(["g++", "-O2", "-I", CSRC, "-o", str(exe), str(src)])
    rng = np.random.default_rng(7)
    cases = ["0", "0.0", "10", "9.98458", ".5", "5.", ".", "..5", "1.2.3", "00010.500000000000000000000",
             "0.1234567890123456789", "123456789012345678", "9007199254740993", "0.30000000000000004",
             "9007199254740994.99999999", "0.1000000000000000055511151231257827021181583404541015625",
             "12345678901234567890", "0." + "0" * 40 + "1", "4.35", "2.675", "1.005", "0.000001"]
    for _ in range(20000):
        ni, nf = int(rng.integers(0, 8)), int(rng.integers(0, 22))
        s = "".join(str(int(x)) for x in rng.integers(0, 10, ni))
        if rng.random() < 0.9:
            s += "." + "".join(str(int(x)) for x in rng.integers(0, 10, nf))
        if s:
            cases.append(s)
    out = subprocess.run([str(exe)], input="\n".join(cases) + "\n", capture_output=True, text=True, check=True).stdout
    lines = out.strip().split("\n")
    assert len(lines) == len(cases)
    unsupported = 0
    for c, ln in zip(cases, lines):
        f = ln.split()
        lit = c
        if c.count(".") > 1:                       # parseFloat stops at the second '.'
            first = c.index(".")
            lit = c[:c.index(".", first + 1)]
        has_digit = any(ch.isdigit() for ch in lit)
        if f[0] == "nan":
            assert not has_digit, c
        elif f[0] == "unsupported":
            unsupported += 1
            assert len(lit.replace(".", "").lstrip("0")) > 19 or "." not in lit or len(lit) > 40, c
        else:
            assert has_digit, c
            assert f[0] == f[1], c                                     # glibc strtod
            assert int(f[0], 16) == int(np.float64(float(lit)).view(np.uint64)), c   # CPython
        assert int(f[-1]) == len(c), c              # the whole [0-9.]+ run is consumed
    assert unsupported < len(cases) // 100      # only >19-digit inputs with an undecidable tail
    assert lines[cases.index("9007199254740994.99999999")].startswith("unsupported")
    assert lines[cases.index("12345678901234567890")].startswith("unsupported")   # 20 integer digits


DIV_HARNESS = r"""
#include "decimal.cuh"
#include <stdio.h>
#include <random>
using namespace hlsmap;
int main() {
    std::mt19937_64 rng(12345);
    long long bad = 0, fell = 0, n = 0;
    auto check = [&](unsigned long long w, int k, bool above) {
        unsigned long long q; int e; bool rem;
        if (!div_pow10_estimate(w, k, &q, &e, &rem)) fell++;
        const double a = div_pow10_rounded<true>(w, k, above), b = div_pow10_rounded<false>(w, k, above);
        n++;
        if (memcmp(&a, &b, 8)) { if (bad < 10) printf("MISMATCH w=%llu k=%d above=%d\n", w, k, (int)above); bad++; }
    };
    for (int k = 0; k <= 19; k++) {
        unsigned long long d = 1; for (int i = 0; i < k; i++) d *= 10ull;
        for (int it = 0; it < 60000; it++) {
            unsigned long long w = rng() >> (rng() % 64);
            if (it % 7 == 0) {            // next to multiples of 10^k: remainders 0 and +-1, ties
                const unsigned long long mult = (rng() >> (rng() % 60)) | 1ull;
                const unsigned __int128 v = (unsigned __int128)mult * d + (long long)(rng() % 5) - 2;
                if (v >> 64) continue;
                w = (unsigned long long)v;
            }
            if (w == 0) continue;
            check(w, k, false);
            check(w, k, true);
        }
        for (int b = 0; b < 64; b++)
            for (int dl = -2; dl <= 2; dl++)
                if ((1ull << b) + dl) check((1ull << b) + dl, k, false);
    }
    printf("%lld %lld %lld\n", n, bad, fell);
    return 0;
}
"""


def test_estimated_division_equals_long_division(tmp_path):
    """decimal.cuh divides by 10^k (k <= 19) with a floating-point estimate corrected exactly in
    integers; the subtract-and-shift division it replaced is kept as the fall-back and is the
    reference here: same bits on ~2.4 million (w, k, sticky) cases, and no case falls back."""
    src, exe = tmp_path / "d.cpp", tmp_path / "d"
    src.write_text(DIV_HARNESS)
    subprocess.check_call(["g++", "-O2", "-I", CSRC, "-o", str(exe), str(src)])
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.strip().split("\n")
    n, bad, fell = (int(x) for x in out[-1].split())
    assert n > 2_000_000 and bad == 0 and fell == 0, out


_random_playlist = synth.m3u8_token_soup


def test_hand_emulation_agrees_with_a_regex_engine():
    """oracle_parse_m3u8 walks the text by hand; oracle.pyref runs the same expression through
    Python's `re`.  Random token soup (hidden tags, cut tags, odd numbers, U+2028) must agree."""
    rng = np.random.default_rng(11)
    texts = [_random_playlist(rng) for _ in range(6000)] + [t.decode() for t in synth._NASTY]
    raw = [t.encode("utf-8") for t in texts]
    off = np.concatenate([[0], np.cumsum([len(b) for b in raw])])
    r = oracle.parse_m3u8(b"".join(raw), off)
    total = 0
    for i, t in enumerate(texts):
        frags, live = pyref.parse_level_playlist(t)
        a, b = int(r["track_off"][i]), int(r["track_off"][i + 1])
        if any(f["sn"] > 2 ** 31 - 1 for f in frags):
            continue                                   # beyond int32: the C side only flags these
        assert b - a == len(frags), repr(t)
        assert [f["sn"] for f in frags] == list(r["sn"][a:b]), repr(t)
        assert [f["duration"] for f in frags] == list(r["duration"][a:b]), repr(t)
        assert [f["start"] for f in frags] == list(r["start"][a:b]), repr(t)
        assert int(live) == r["live"][i], repr(t)
        total += len(frags)
    assert total > 800


def test_text_shards_parse_to_the_slices_of_the_whole():
    from hlsjs_p2p_wrapper_b200 import shard
    base, state, texts = synth.m3u8_corpus(n_playlists=90, unparsed_every=13)
    off = np.concatenate([[0], np.cumsum([len(t) for t in texts])]).astype(np.int64)
    text = b"".join(texts)
    whole = oracle.parse_m3u8(text, off, state)
    bounds = np.asarray([0, 17, 17, 60, 90])                     # one empty shard on purpose
    parts = shard.split_texts(base, state, text, off, bounds)
    f0 = 0
    for (b, st, tx, to), p0, p1 in zip(parts, bounds[:-1], bounds[1:]):
        assert b[0] == 0 and b[-1] == len(st) == len(to) - 1 and to[0] == 0 and to[-1] == len(tx)
        r = oracle.parse_m3u8(tx, to, st)
        n = int(r["track_off"][-1])
        assert np.array_equal(r["sn"], whole["sn"][f0:f0 + n])
        assert np.array_equal(r["start"], whole["start"][f0:f0 + n])
        t0, t1 = base[p0], base[p1]
        assert np.array_equal(r["track_off"], whole["track_off"][t0:t1 + 1] - whole["track_off"][t0])
        f0 += n
    assert f0 == whole["track_off"][-1]

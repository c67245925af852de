"""CPU-side checks of the JSON form of a SegmentView: the oracle's Number::toString
against CPython's repr digits (a third, unrelated implementation), the reference's own toJSON test
(test/segment-view.js:5-11, in tests/golden/reference_vectors.json), and the product's shortest-digit
generator (csrc/dtoa.cuh, compiled for the host) against both."""
import decimal
import json
import os
import struct
import subprocess

import numpy as np

import oracle
from oracle import pyref
from hlsjs_p2p_wrapper_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "hlsjs-p2p-wrapper_b200", "csrc")


def js_from_repr(x):
    """Number::toString laid out from CPython's shortest round-trip digits."""
    if x != x:
        return "NaN"
    if x == 0:
        return "0"
    if x in (float("inf"), float("-inf")):
        return "Infinity" if x > 0 else "-Infinity"
    sign, digs, exp = decimal.Decimal(repr(abs(x))).as_tuple()
    digs = list(digs)
    while len(digs) > 1 and digs[-1] == 0:
        digs.pop()
        exp += 1
    while len(digs) > 1 and digs[0] == 0:
        digs.pop(0)
    k, n = len(digs), exp + len(digs)
    d = "".join(map(str, digs))
    if k <= n <= 21:
        s = d + "0" * (n - k)
    elif 0 < n <= 21:
        s = d[:n] + "." + d[n:]
    elif -6 < n <= 0:
        s = "0." + "0" * (-n) + d
    else:
        s = d[0] + ("." + d[1:] if k > 1 else "") + "e" + ("-" if n - 1 < 0 else "+") + str(abs(n - 1))
    return ("-" if x < 0 else "") + s


_doubles = synth.sample_doubles


def test_oracle_number_to_string_agrees_with_cpython_repr():
    for x in _doubles(40000, 1):
        assert oracle.js_number_to_string(float(x)) == js_from_repr(float(x)), repr(x)
    assert oracle.js_number_to_string(2.0 ** -24) == "5.960464477539063e-8"      # 16 digits, not 17
    assert oracle.js_number_to_string(1e21) == "1e+21" and oracle.js_number_to_string(1e20) == "1" + "0" * 20
    assert oracle.js_number_to_string(-0.0) == "0" and oracle.js_number_to_string(1e-7) == "1e-7"


def test_oracle_json_matches_python_json_and_the_reference_test(golden):
    rng = np.random.default_rng(2)
    n = 3000
    lv, ui, sn = (rng.integers(-5, 50, n), rng.integers(0, 4, n), rng.integers(-2 ** 31, 2 ** 31, n))
    tm = _doubles(4 * n, 3)[:n].copy()
    tm[::97] = np.nan
    tm[5::211] = np.inf
    off, chars = oracle.view_to_json(lv, ui, sn, tm)
    text = bytes(chars)
    for i in range(n):
        obj = json.loads(text[off[i]:off[i + 1]])
        assert list(obj) == ["sn", "trackView", "time"] and list(obj["trackView"]) == ["level", "urlId"]
        assert (obj["sn"], obj["trackView"]["level"], obj["trackView"]["urlId"]) == (sn[i], lv[i], ui[i])
        if np.isfinite(tm[i]):
            assert obj["time"] == tm[i] and (obj["time"] != 0 or True)
        else:
            assert obj["time"] is None
    back = oracle.view_from_json(text, off)
    assert not back["status"].any()
    assert np.array_equal(back["sn"], sn) and np.array_equal(back["level"], lv) and np.array_equal(back["url_id"], ui)
    fin = np.isfinite(tm)
    assert np.array_equal(back["time"][fin], tm[fin]) and np.isnan(back["time"][~fin]).all()
    assert ((back["flags"] & 32) != 0).tolist() == (~fin).tolist()
    # the reference's own test: a view with every field undefined survives the trip
    for v in golden["segmentView"]["json_roundtrip"]:
        sv = pyref.SegmentView(v["ctor"])
        s = sv.json_stringify()
        assert s == v["derived_json"]
        r = oracle.view_from_json(s.encode(), [0, len(s)])
        assert r["status"][0] == 0 and r["flags"][0] == 2          # only trackView, and it is empty
    # the mirror's stringify of a complete view is the batch text
    sv = pyref.SegmentView({"sn": 25, "trackView": {"level": 1, "urlId": 0}, "time": 12.5})
    o, c = oracle.view_to_json([1], [0], [25], [12.5])
    assert bytes(c).decode() == sv.json_stringify() == '{"sn":25,"trackView":{"level":1,"urlId":0},"time":12.5}'


def test_oracle_json_parse_rejects_other_shapes():
    bad = [b'', b'{', b'{"sn":1', b'{"sn":01,"trackView":{}}', b'{"sn":1,"trackView":{"urlId":1,"level":2}}',
           b'{"sn": 1,"trackView":{}}', b'{"sn":1,"trackView":{}} ', b'{"time":nul}', b'{"sn":1.5,"trackView":{}}',
           b'{"sn":3000000000,"trackView":{}}', b'{"sn":1,"trackView":{},"time":1e}', b'[1]']
    off = np.concatenate([[0], np.cumsum([len(b) for b in bad])]).astype(np.int32)
    r = oracle.view_from_json(b"".join(bad), off)
    assert list(r["status"]) == [1, 1, 1, 1, 1, 1, 1, 1, 3, 3, 1, 1]
    ok = [b'{}', b'{"trackView":{}}', b'{"time":1e3}', b'{"sn":-0,"trackView":{"urlId":2}}', b'{"sn":2.5e1,"time":null}']
    off = np.concatenate([[0], np.cumsum([len(b) for b in ok])]).astype(np.int32)
    r = oracle.view_from_json(b"".join(ok), off)
    assert not r["status"].any() and list(r["flags"]) == [0, 2, 16, 1 | 2 | 8, 1 | 16 | 32]
    assert r["time"][2] == 1000.0 and r["sn"][4] == 25


HARNESS = r"""
#include "dtoa.cuh"
#include <stdio.h>
using namespace hlsmap;
static const unsigned long long T[][2] = {
#include "pow10_table.inc"
};
int main() {
    unsigned long long b;
    while (scanf("%llx", &b) == 1) {
        Shortest s = shortest_digits(b & 0x7fffffffffffffffull, T);
        char out[64];
        int n = js_number_write(s, out);
        out[n] = 0;
        printf("%s %d\n", out, js_number_length(s));
    }
}
"""


def test_device_digit_generator_on_the_host(tmp_path):
    src, exe = tmp_path / "h.cpp", tmp_path / "h"
    src.write_text(HARNESS)
    subprocess.check_call(["g++", "-O2", "-I", CSRC, "-o", str(exe), str(src)])
    x = _doubles(200000, 5)
    x = np.abs(x[x != 0])
    inp = "\n".join("%x" % struct.unpack("<Q", struct.pack("<d", float(v)))[0] for v in x) + "\n"
    out = subprocess.run([str(exe)], input=inp, capture_output=True, text=True, check=True).stdout.split("\n")
    assert len(out) >= len(x)
    for v, ln in zip(x, out):
        s, n = ln.split()
        assert s == js_from_repr(float(v)), repr(v)
        assert int(n) == len(s)
    for v in x[:20000]:
        assert oracle.js_number_to_string(float(v)) == js_from_repr(float(v))


SWAR_HARNESS = r"""
#include "swar.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
using namespace hlsmap;
int main() {
    unsigned long long bad = 0;
    srand(7);
    for (long it = 0; it < 3000000; it++) {
        char s[8];
        const int nd = rand() % 9;
        for (int i = 0; i < 8; i++) {
            if (i < nd) s[i] = (char)('0' + rand() % 10);
            else { int c; do { c = rand() % 256; } while (i == nd && c >= '0' && c <= '9'); s[i] = (char)c; }
        }
        unsigned long long v;
        memcpy(&v, s, 8);
        int exp = 0;
        while (exp < 8 && s[exp] >= '0' && s[exp] <= '9') exp++;
        if (leading_digits8(v) != exp) bad++;
        for (int k = 1; k <= exp; k++) {
            unsigned e = 0;
            for (int i = 0; i < k; i++) e = e * 10 + (unsigned)(s[i] - '0');
            if (digits8_value(v, k) != e) bad++;
        }
    }
    printf("%llu\n", bad);
    return bad != 0;
}
"""


def test_word_at_a_time_digit_routines_on_the_host(tmp_path):
    """csrc/swar.cuh (what the JSON reader takes its int32 columns with): the count of leading digit
    bytes and the value of 1..8 digits of a little-endian word, against byte loops on random words."""
    src, exe = tmp_path / "s.cpp", tmp_path / "s"
    src.write_text(SWAR_HARNESS)
    subprocess.check_call(["g++", "-O2", "-I", CSRC, "-o", str(exe), str(src)])
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.strip()
    assert out == "0"


READERS_HARNESS = r"""
#include "swar.cuh"
#include <cstdio>
#include <cstring>
#include <string>
#include <iostream>
using namespace hlsmap;
static const double P10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                               1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
static const unsigned long long IP10[20] = {1ull, 10ull, 100ull, 1000ull, 10000ull, 100000ull, 1000000ull, 10000000ull,
    100000000ull, 1000000000ull, 10000000000ull, 100000000000ull, 1000000000000ull, 10000000000000ull,
    100000000000000ull, 1000000000000000ull, 10000000000000000ull, 100000000000000000ull, 1000000000000000000ull,
    10000000000000000000ull};
struct Get8 {
    const char *s;          // zero-padded copy of the text
    unsigned long long operator()(long long p) const { unsigned long long v; memcpy(&v, s + p, 8); return v; }
};
int main() {
    std::string line;
    while (std::getline(std::cin, line)) {           // "<i|d> <text>"
        const char kind = line[0];
        std::string txt = line.substr(2);
        const long long lim = (long long)txt.size();
        std::string padded = txt + std::string(32, '7');     // garbage digits beyond lim must not count
        Get8 g{padded.c_str()};
        long long q = 0;
        if (kind == 'i') {
            int v = 0;
            const bool ok = swar_int32(g, q, lim, &v);
            printf("%d %d %lld\n", ok ? 1 : 0, v, q);
        } else {
            double x = 0;
            const bool ok = swar_plain_decimal(g, q, lim, &x, P10, IP10);
            unsigned long long b;
            memcpy(&b, &x, 8);
            printf("%d %llx %lld\n", ok ? 1 : 0, b, q);
        }
    }
}
"""


def test_word_at_a_time_readers_on_the_host(tmp_path):
    """swar_int32 / swar_plain_decimal (the JSON reader's fast path) against Python's int() and
    float(): whenever they accept a text they must give exactly that value and consume exactly the
    number; shapes they do not cover must be declined, not guessed."""
    import random
    src, exe = tmp_path / "r.cpp", tmp_path / "r"
    src.write_text(READERS_HARNESS)
    subprocess.check_call(["g++", "-O2", "-I", CSRC, "-o", str(exe), str(src)])
    rnd = random.Random(11)
    cases = []
    ints = [0, 1, -1, 9, 10, 99999999, 100000000, 123456789, 2147483647, -2147483648, 2147483648, -2147483649,
            4294967296, 99999999999]
    ints += [rnd.randrange(-2 ** 31, 2 ** 31) for _ in range(20000)]
    ints += [rnd.randrange(-10 ** k, 10 ** k) for k in range(1, 11) for _ in range(300)]
    for v in ints:
        for tail in ("", ",", "}", "x"):
            cases.append(("i", str(v) + tail, v, len(str(v))))
    for bad in ("", "-", "01", "-01", "1.5", "1e3", "1E3", "12345678901", "a1", "--1", "+1"):
        cases.append(("i", bad, None, 0))
    decs = ["0", "0.5", "12.5", "86399.999", "0.1", "0.30000000000000004", "1234567.123456789012", "9999999.99999999",
            "0.000001", "5e-7", "1.", ".5", "00.5", "12345678.5", "1.1234567890123456", "0.1234567890123456",
            "86399.99900000001", "-0.0", "-12.75", "1.5e3", "7", "1234567"]
    for _ in range(40000):
        ni, nf = rnd.randrange(1, 8), rnd.randrange(0, 17)
        ip = str(rnd.randrange(10 ** (ni - 1) if ni > 1 else 0, 10 ** ni))
        fp = "".join(rnd.choice("0123456789") for _ in range(nf))
        decs.append(("-" if rnd.random() < 0.1 else "") + ip + ("." + fp if nf else ""))
    for _ in range(20000):                  # what JSON.stringify prints for playlist times
        decs.append(repr(rnd.randrange(0, 86400000) / 1000.0 + rnd.choice([0.0, 1e-9, 2.5e-7])))
    for d in decs:
        for tail in ("", "}", ",\""):
            cases.append(("d", d + tail, d, len(d)))
    inp = "".join("%s %s\n" % (k, t) for k, t, _, _ in cases)
    out = subprocess.run([str(exe)], input=inp, capture_output=True, text=True, check=True).stdout.split("\n")
    accepted = {"i": 0, "d": 0}
    for (kind, text, want, n), ln in zip(cases, out):
        ok, val, q = ln.split()
        if ok == "0":
            continue
        accepted[kind] += 1
        assert int(q) == n, (text, ln)
        if kind == "i":
            assert want is not None and -2 ** 31 <= want < 2 ** 31 and int(val) == want, (text, ln)
        else:
            f = float(want)          # accepted => plain decimal => Python reads the same value
            assert "e" not in want and "E" not in want, text
            assert struct.unpack("<Q", struct.pack("<d", f))[0] == int(val, 16), (text, ln)
    # the shapes the fast path exists for are accepted, not silently pushed to the slow path
    assert accepted["i"] > 4 * 20000 and accepted["d"] > 3 * 45000

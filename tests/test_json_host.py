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

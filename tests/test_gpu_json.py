"""JSON.stringify(segmentView) / new SegmentView(JSON.parse(text)) in batch on the GPU (SURVEY 8 row
f3; test/segment-view.js:5-11) against oracle/: byte-identical text, bit-identical columns back."""
import json

import numpy as np
import pytest

import oracle
from hlsjs_p2p_wrapper_b200 import synth, _native as N

_doubles = synth.sample_doubles

pytestmark = pytest.mark.gpu


def _views(n, seed):
    rng = np.random.default_rng(seed)
    return (rng.integers(-3, 40, n).astype(np.int32), rng.integers(0, 5, n).astype(np.int32),
            rng.integers(-2 ** 31, 2 ** 31, n).astype(np.int32))


def test_stringify_is_byte_identical(ctx):
    n = 100_000
    lv, ui, sn = _views(n, 1)
    tm = np.resize(_doubles(60000, 2), n).copy()
    tm[::101] = np.nan
    tm[7::503] = -np.inf
    tm[11::89] = 0.0
    tm[13::97] = -0.0
    off, chars = ctx.view_to_json(lv, ui, sn, tm)
    eoff, echars = oracle.view_to_json(lv, ui, sn, tm)
    assert np.array_equal(off, eoff) and np.array_equal(chars, echars)
    # and it is JSON: Python's own parser reads the same values back
    text = bytes(chars)
    for i in range(0, n, 997):
        o = json.loads(text[off[i]:off[i + 1]])
        assert o["sn"] == sn[i] and o["trackView"] == {"level": lv[i], "urlId": ui[i]}
        assert (o["time"] is None) if not np.isfinite(tm[i]) else o["time"] == tm[i]
    # a view without time (fromArrayBuffer builds those): the property is left out
    off, chars = ctx.view_to_json(lv[:1000], ui[:1000], sn[:1000])
    eoff, echars = oracle.view_to_json(lv[:1000], ui[:1000], sn[:1000])
    assert np.array_equal(off, eoff) and np.array_equal(chars, echars)
    assert bytes(chars[:off[1]]).decode() == '{"sn":%d,"trackView":{"level":%d,"urlId":%d}}' % (sn[0], lv[0], ui[0])


def test_times_of_a_playlist_print_like_the_reference_would(ctx):
    t, q = synth.config3(n_players=2000)
    res = ctx.upload(t).get_segment_list(q.playlist, q.level, q.begin, q.duration)
    n = len(res.sn)
    lv = np.repeat(q.level, np.diff(res.offsets)).astype(np.int32)
    off, chars = ctx.view_to_json(lv, np.zeros(n, np.int32), res.sn, res.time)
    eoff, echars = oracle.view_to_json(lv, np.zeros(n, np.int32), res.sn, res.time)
    assert n > 5000 and np.array_equal(off, eoff) and np.array_equal(chars, echars)


def test_parse_round_trip_and_status(ctx):
    n = 50_000
    lv, ui, sn = _views(n, 3)
    tm = np.resize(_doubles(40000, 4), n).copy()
    tm = np.where((np.abs(tm) < 1e19) & ((np.abs(tm) > 1e-20) | (tm == 0)), tm, 12.5)   # the converter's range
    tm[::77] = np.nan
    off, chars = ctx.view_to_json(lv, ui, sn, tm)
    got = ctx.view_from_json(chars, off)
    exp = oracle.view_from_json(chars, off)
    for k in ("status", "flags", "sn", "level", "url_id"):
        assert np.array_equal(got[k], exp[k]), k
    assert np.array_equal(got["time"].view(np.int64), exp["time"].view(np.int64))
    assert not got["status"].any()
    fin = np.isfinite(tm)
    assert np.array_equal(got["time"][fin], tm[fin]) and np.array_equal(got["sn"], sn)
    # isEqual of the transferred view, as the reference's test asserts (segment-view.js:33-39)
    assert ctx.view_is_equal(lv, ui, sn, got["level"], got["url_id"], got["sn"]).all()


def test_parse_other_shapes(ctx):
    texts = [b'', b'{', b'{"sn":1', b'{"sn":01,"trackView":{}}', b'{"sn":1,"trackView":{"urlId":1,"level":2}}',
             b'{"sn": 1,"trackView":{}}', b'{"sn":1,"trackView":{}} ', b'{"time":nul}', b'{"sn":1.5,"trackView":{}}',
             b'{"sn":3000000000,"trackView":{}}', b'{"sn":1,"trackView":{},"time":1e}', b'[1]',
             b'{}', b'{"trackView":{}}', b'{"time":1e3}', b'{"sn":-0,"trackView":{"urlId":2}}',
             b'{"sn":2.5e1,"time":null}', b'{"sn":7,"trackView":{"level":1,"urlId":0},"time":0.30000000000000004}',
             b'{"time":-1.5E-7}', b'{"time":100000000000000000000}']
    off = np.concatenate([[0], np.cumsum([len(b) for b in texts])]).astype(np.int32)
    got = ctx.view_from_json(b"".join(texts), off)
    exp = oracle.view_from_json(b"".join(texts), off)
    for k in ("status", "flags", "sn", "level", "url_id"):
        assert np.array_equal(got[k], exp[k]), k
    assert np.array_equal(got["time"], exp["time"], equal_nan=True)
    assert got["time"][-3] == 0.30000000000000004 and got["time"][-2] == -1.5e-7 and got["time"][-1] == 1e20
    # what the exact converter does not cover is reported, not guessed
    far = [b'{"time":1e300}', b'{"time":5e-324}', b'{"sn":1e40}', b'{"time":123456789012345678901234}']
    off = np.concatenate([[0], np.cumsum([len(b) for b in far])]).astype(np.int32)
    got = ctx.view_from_json(b"".join(far), off)
    assert list(got["status"]) == [N.JSON_UNSUPPORTED] * 4


def test_capacity_and_empty(ctx):
    off, chars = ctx.view_to_json([], [], [], [])
    assert list(off) == [0] and len(chars) == 0
    import ctypes
    lv, ui, sn = _views(10, 5)
    offsets, buf, total = np.zeros(11, np.int32), np.empty(20, np.uint8), ctypes.c_int64(0)
    p = lambda a: ctypes.c_void_p(a.ctypes.data)
    rc = N.load().hlsmap_view_to_json(ctx._h, 10, p(lv), p(ui), p(sn), None, p(offsets), p(buf), 20, ctypes.byref(total))
    assert rc == N.ERR_CAPACITY and total.value == offsets[10] > 20


def test_async_calls_of_growing_size_on_a_caller_stream(ctx):
    """Two device-pointer calls queued back to back on the caller's stream, the second needing larger
    scratch than the first: replacing the scratch must not pull it from under the first call."""
    import ctypes
    import torch
    lib, dev = N.load(), torch.device("cuda", 0)
    stream = torch.cuda.Stream()
    v = lambda x: ctypes.c_void_p(x.data_ptr())
    outs = []
    torch.cuda.synchronize()
    for n in (300_000, 3_000_000):
        lv, ui, sn = _views(n, n)
        tm = np.resize(_doubles(50000, 9), n).copy()
        d = [torch.from_numpy(a).to(dev) for a in (lv, ui, sn, tm)]
        d_off = torch.empty(n + 1, dtype=torch.int32, device=dev)
        d_ch = torch.empty(104 * n, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        outs.append((lv, ui, sn, tm, d, d_off, d_ch))
    for lv, ui, sn, tm, d, d_off, d_ch in outs:                      # both queued before either is awaited
        N.check(lib.hlsmap_view_to_json_dev(ctx._h, len(sn), v(d[0]), v(d[1]), v(d[2]), v(d[3]), v(d_off), v(d_ch),
                                            104 * len(sn), ctypes.c_void_p(stream.cuda_stream), None))
    torch.cuda.synchronize()
    for lv, ui, sn, tm, d, d_off, d_ch in outs:
        eoff, echars = oracle.view_to_json(lv[:20000], ui[:20000], sn[:20000], tm[:20000])
        off = d_off.cpu().numpy()
        assert np.array_equal(off[:20001], eoff)
        assert np.array_equal(d_ch[:off[20000]].cpu().numpy(), echars)

"""Level playlists (M3U8 text) -> fragment tables on the GPU (hlsmap_parse_m3u8),
against the CPU restatement of hls.js 0.5.46's parseLevelPlaylist in oracle/ (glibc strtod is the
independent decimal conversion).  Bit-exact: `start` is compared as int64 bit patterns."""
import ctypes

import numpy as np
import pytest

import oracle
from hlsjs_p2p_wrapper_b200 import synth, _native as N
from hlsjs_p2p_wrapper_b200.tables import FragmentTables, TRACK_PARSED

pytestmark = pytest.mark.gpu


def _offsets(texts):
    return np.concatenate([[0], np.cumsum([len(t) for t in texts])]).astype(np.int64)


def _same(got, exp):
    assert np.array_equal(got["track_off"], exp["track_off"])
    assert np.array_equal(got["live"], exp["live"])
    assert np.array_equal(got["sn"], exp["sn"])
    assert np.array_equal(got["duration"].view(np.int64), exp["duration"].view(np.int64))
    assert np.array_equal(got["start"].view(np.int64), exp["start"].view(np.int64))


def test_one_playlist_spelled_out(ctx):
    text = (b"#EXTM3U\n#EXT-X-TARGETDURATION:10\n#EXT-X-MEDIA-SEQUENCE:2680\n#EXTINF:9.9,\nseg2680.ts\n"
            b"#EXTINF:10,\nseg2681.ts\n#EXTINF:0.1,\nseg2682.ts\n")
    r = ctx.parse_m3u8([text])
    assert list(r["sn"]) == [2680, 2681, 2682] and list(r["live"]) == [1]
    assert list(r["duration"]) == [9.9, 10.0, 0.1]
    assert list(r["start"]) == [0.0, 9.9, 9.9 + 10.0]          # the running sum, in order
    assert r["result"].n_fragments == 3 and r["result"].n_sequence_tags == 1
    r = ctx.parse_m3u8([text + b"#EXT-X-ENDLIST\n"])
    assert list(r["live"]) == [0]


def test_corner_cases_of_the_regular_expression(ctx):
    for k, text in enumerate(synth._NASTY):
        got = ctx.parse_m3u8([text])
        exp = oracle.parse_m3u8(text, [0, len(text)])
        assert not exp["bad"], k
        _same(got, exp)


def test_corpus_with_ignored_tracks(ctx):
    base, state, texts = synth.m3u8_corpus(n_playlists=400, unparsed_every=17)
    off = _offsets(texts)
    got = ctx.parse_m3u8(texts, track_state=state)
    exp = oracle.parse_m3u8(b"".join(texts), off, state)
    _same(got, exp)
    assert got["result"].n_fragments == exp["track_off"][-1] > 5000
    lens = np.diff(got["track_off"])
    assert not lens[state != TRACK_PARSED].any()


def test_tags_across_tile_borders(ctx):
    """Every alignment of a playlist against a tile border (16 KB tiles; 32 KB is every second one),
    including tags and numbers cut by it."""
    body = b"#EXT-X-MEDIA-SEQUENCE:41\n#EXTINF:9.98458,\nfileSequence41.ts\n#EXTINF:3.003,\nb.ts\n#EXT-X-ENDLIST\n"
    for pad in list(range(16384 - len(body) - 2, 16384 + 2)) + list(range(32768 - len(body) - 2, 32768 + 2, 3)):
        texts = [b"x" * pad, body, b"\n", body]
        got = ctx.parse_m3u8(texts)
        exp = oracle.parse_m3u8(b"".join(texts), _offsets(texts))
        _same(got, exp)
        assert list(got["sn"]) == [41, 42, 41, 42]


def test_tile_dense_with_hashes_takes_the_four_round_path(ctx):
    rng = np.random.default_rng(5)
    parts = []
    for i in range(30000):
        parts.append(b"#EXTINF:%d.%d,\nu\n" % (rng.integers(1, 9), rng.integers(0, 999)) if i % 6 == 0 else b"#\n")
    text = b"#EXT-X-MEDIA-SEQUENCE:9\n" + b"".join(parts)
    cut = [0, 1000, 1000, 30001, 70003, len(text)]          # tracks cut the text anywhere, one empty
    texts = [text[a:b] for a, b in zip(cut[:-1], cut[1:])]
    got = ctx.parse_m3u8(texts)
    exp = oracle.parse_m3u8(text, np.asarray(cut, np.int64))
    _same(got, exp)
    assert got["result"].n_fragments > 4000


def test_long_tracks_and_sequence_tags_in_the_middle(ctx):
    rng = np.random.default_rng(6)
    texts = []
    for t in range(40):
        lines = []
        for i in range(int(rng.integers(0, 3000))):
            if rng.random() < 0.004:
                lines.append(b"#EXT-X-MEDIA-SEQUENCE:%d" % rng.integers(0, 10 ** 9))
            lines.append(b"#EXTINF:%.5f,\nseg%d.ts" % (rng.uniform(0.5, 12.0), i))
        texts.append(b"\n".join(lines))
    got = ctx.parse_m3u8(texts)
    exp = oracle.parse_m3u8(b"".join(texts), _offsets(texts), threads=4)
    _same(got, exp)
    assert got["result"].n_sequence_tags > 50


def test_large_regular_block(ctx):
    text, off = synth.m3u8_block(n_tracks=3000, n_frag=400)            # 75 MB, 1.2e6 fragments
    got = ctx.parse_m3u8(text, off)
    exp = oracle.parse_m3u8(text, off, threads=8)
    _same(got, exp)
    base, state, texts = synth.m3u8_corpus(n_playlists=30000, n_levels=3, frags=(6, 6), nasty=False)
    got = ctx.parse_m3u8(texts, track_state=state)                      # 9e4 live windows of 6
    exp = oracle.parse_m3u8(b"".join(texts), _offsets(texts), state, threads=8)
    _same(got, exp)


def test_numbers_that_cannot_be_rounded_exactly_fail_loudly(ctx):
    ok = b"#EXTINF:0.1000000000000000055511151231257827021181583404541015625,\n"   # == the double 0.1
    r = ctx.parse_m3u8([ok])
    assert r["duration"][0] == 0.1
    # 24 digits just below the midpoint of two doubles: the 19 digits kept cannot tell which side
    bad = b"#EXTINF:9007199254740994.99999999,\n"
    with pytest.raises(N.HlsMapError) as e:
        ctx.parse_m3u8([bad])
    assert e.value.code == N.ERR_UNSUPPORTED
    with pytest.raises(N.HlsMapError) as e:
        ctx.parse_m3u8([b"#EXT-X-MEDIA-SEQUENCE:2147483647\n#EXTINF:1,\n#EXTINF:1,\n"])
    assert e.value.code == N.ERR_UNSUPPORTED
    with pytest.raises(N.HlsMapError) as e:
        ctx.parse_m3u8([b"#EXT-X-MEDIA-SEQUENCE:99999999999\n#EXTINF:1,\n"])
    assert e.value.code == N.ERR_UNSUPPORTED


def test_tags_far_into_one_line_are_bounded_work(ctx):
    """Whether an earlier tag on the same line hides a '#' is decided by replaying the line, one
    thread per '#': quadratic in the line's length.  Lines up to 16 KB are replayed (and agree with the
    oracle); a tag further into its line makes the call fail instead of stalling the GPU on a
    megabyte-long line of tags from the network."""
    import time
    line = b"#EXT-X-KEY:" + b"x" * 9000 + b"#EXTINF:5,#EXTINF:6,\n#EXTINF:7,\nseg.ts\n"     # hidden ones: 9 KB in
    got = ctx.parse_m3u8([line])
    exp = oracle.parse_m3u8(line, [0, len(line)])
    _same(got, exp)
    assert list(got["duration"]) == [7.0]
    hostile = b"#EXTINF:1," * 250_000 + b"\n#EXTINF:4,\nseg.ts\n"          # one 2.5 MB line of tags
    t0 = time.perf_counter()
    with pytest.raises(N.HlsMapError) as e:
        ctx.parse_m3u8([hostile])
    assert e.value.code == N.ERR_UNSUPPORTED and "16384" in e.value.message
    assert time.perf_counter() - t0 < 5.0


def test_capacity_protocol_and_empty_inputs(ctx):
    lib = N.load()
    text = b"#EXTINF:1,\n#EXTINF:2,\n#EXTINF:3,\n"
    buf = np.frombuffer(text, np.uint8)
    off = np.asarray([0, len(text)], np.int64)
    toff, live = np.zeros(2, np.int64), np.zeros(1, np.uint8)
    st, sn, du = np.empty(2), np.empty(2, np.int32), np.empty(2)
    res = N.IngestResult()
    p = lambda a: ctypes.c_void_p(a.ctypes.data)
    rc = lib.hlsmap_parse_m3u8(ctx._h, 1, None, p(buf), p(off), p(toff), p(live), p(st), p(sn), p(du), 2,
                               ctypes.byref(res))
    assert rc == N.ERR_CAPACITY and res.n_fragments == 3 and list(toff) == [0, 3]
    r = ctx.parse_m3u8([])
    assert list(r["track_off"]) == [0] and r["result"].n_fragments == 0
    r = ctx.parse_m3u8([b"", b"", b""])
    assert list(r["track_off"]) == [0, 0, 0, 0] and list(r["live"]) == [1, 1, 1]


def test_map_from_text_answers_like_map_from_tables(ctx):
    base, state, texts = synth.m3u8_corpus(n_playlists=300, unparsed_every=19)
    exp = oracle.parse_m3u8(b"".join(texts), _offsets(texts), state)
    tables = FragmentTables(base, exp["track_off"], state, np.ones(len(state), np.int32), exp["start"],
                            exp["sn"], exp["duration"]).validate()
    m_text = ctx.upload_m3u8(base, state, texts)
    m_tab = ctx.upload(tables)
    assert m_text.ingest.n_fragments == tables.n_fragments
    a, b = m_text.info(), m_tab.info()
    for k in ("n_fragments", "n_parsed_tracks", "n_monotone_tracks", "n_tracks", "n_playlists"):
        assert a[k] == b[k], k
    rng = np.random.default_rng(3)
    n = 20000
    pl = rng.integers(0, 300, n).astype(np.int32)
    lv = rng.integers(0, 3, n).astype(np.int32)
    bg = rng.uniform(-5, 80, n)
    du = rng.uniform(0, 40, n)
    ra, rb = m_text.get_segment_list(pl, lv, bg, du), m_tab.get_segment_list(pl, lv, bg, du)
    assert np.array_equal(ra.offsets, rb.offsets) and np.array_equal(ra.sn, rb.sn)
    assert np.array_equal(ra.time, rb.time) and np.array_equal(ra.status, rb.status)
    da, sa = m_text.get_segment_duration(pl, lv)
    db, sb = m_tab.get_segment_duration(pl, lv)
    assert np.array_equal(da, db, equal_nan=True) and np.array_equal(sa, sb)
    m_text.close()
    m_tab.close()


def test_device_pointer_entry(ctx):
    import torch
    text, off = synth.m3u8_block(n_tracks=500, n_frag=50)
    dev = torch.device("cuda", 0)
    d_text = torch.from_numpy(text.copy()).to(dev)
    d_off = torch.from_numpy(off).to(dev)
    t = len(off) - 1
    d_toff = torch.empty(t + 1, dtype=torch.int64, device=dev)
    d_live = torch.empty(t, dtype=torch.uint8, device=dev)
    cap = t * 50
    d_start = torch.empty(cap, dtype=torch.float64, device=dev)
    d_sn = torch.empty(cap, dtype=torch.int32, device=dev)
    d_dur = torch.empty(cap, dtype=torch.float64, device=dev)
    res = N.IngestResult()
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()
    v = lambda x: ctypes.c_void_p(x.data_ptr())
    N.check(N.load().hlsmap_parse_m3u8_dev(ctx._h, t, None, v(d_text), v(d_off), len(text), v(d_toff), v(d_live),
                                           v(d_start), v(d_sn), v(d_dur), cap, ctypes.byref(res),
                                           ctypes.c_void_p(stream.cuda_stream)))
    exp = oracle.parse_m3u8(text, off)
    got = dict(track_off=d_toff.cpu().numpy(), live=d_live.cpu().numpy(), start=d_start.cpu().numpy(),
               sn=d_sn.cpu().numpy(), duration=d_dur.cpu().numpy())
    _same(got, exp)


def test_token_soup(ctx):
    """Text glued at random from the tokens the expression cares about (the same soup the CPU suite
    feeds to a real regex engine, tests/test_ingest_host.py)."""
    rng = np.random.default_rng(11)
    import re
    texts = [synth.m3u8_token_soup(rng).encode() for _ in range(30000)]
    # numbers glued into more than 15 characters are the exact converter's "unsupported" territory
    # (and sequence numbers beyond int32): those inputs have their own test above
    texts = [t for t in texts if not re.search(rb"[0-9.]{16,}", t) and not re.search(rb"SEQUENCE:[0-9]{10,}", t)]
    assert len(texts) > 8000
    got = ctx.parse_m3u8(texts)
    exp = oracle.parse_m3u8(b"".join(texts), _offsets(texts))
    assert not exp["bad"]
    _same(got, exp)
    assert got["result"].n_fragments > 800


def test_sharded_map_from_text(ctx):
    """Two shards (two contexts on this GPU), each parsing its own playlists; answers must be those
    of the single map."""
    from hlsjs_p2p_wrapper_b200 import shard
    from hlsjs_p2p_wrapper_b200.engine import Context
    from hlsjs_p2p_wrapper_b200.tables import QueryBatch
    base, state, texts = synth.m3u8_corpus(n_playlists=120, unparsed_every=23)
    off = _offsets(texts)
    text = b"".join(texts)
    other = Context(0)
    sm = shard.ShardedMap.from_m3u8([ctx, other], base, state, text, off)
    one = ctx.upload_m3u8(base, state, texts)
    rng = np.random.default_rng(4)
    n = 8000
    q = QueryBatch(rng.integers(-1, 121, n).astype(np.int32), rng.integers(0, 4, n).astype(np.int32),
                   rng.uniform(-5, 80, n), rng.uniform(0, 40, n))
    o, st, sn, tm = sm.get_segment_list(q)
    r = one.get_segment_list(q.playlist, q.level, q.begin, q.duration)
    assert np.array_equal(o, r.offsets.astype(np.int64)) and np.array_equal(st, r.status)
    assert np.array_equal(sn, r.sn) and np.array_equal(tm, r.time)
    assert 0 < sm.bounds[1] < 120 and len(sn) > 1000
    sm.close()
    one.close()
    other.close()


def test_map_from_text_already_on_the_device(ctx):
    import ctypes
    import torch
    from hlsjs_p2p_wrapper_b200.engine import DeviceMap
    base, state, texts = synth.m3u8_corpus(n_playlists=150, unparsed_every=29)
    off = _offsets(texts)
    text = np.frombuffer(b"".join(texts), np.uint8)
    dev = torch.device("cuda", 0)
    d_text, d_off = torch.from_numpy(text.copy()).to(dev), torch.from_numpy(off).to(dev)
    torch.cuda.synchronize()
    m = object.__new__(DeviceMap)
    m.ctx, m._lib, m.n_playlists = ctx, ctx._lib, len(base) - 1
    h, res = ctypes.c_void_p(), N.IngestResult()
    p = lambda a: ctypes.c_void_p(a.ctypes.data)
    N.check(N.load().hlsmap_map_create_from_m3u8_dev(ctx._h, len(base) - 1, p(base), len(state), p(state), None,
                                                     ctypes.c_void_p(d_text.data_ptr()), ctypes.c_void_p(d_off.data_ptr()),
                                                     len(text), ctypes.byref(res), ctypes.byref(h)))
    m._h = h
    one = ctx.upload_m3u8(base, state, texts)
    assert res.n_fragments == one.ingest.n_fragments > 1000
    rng = np.random.default_rng(8)
    n = 10000
    pl, lv = rng.integers(0, 150, n).astype(np.int32), rng.integers(0, 3, n).astype(np.int32)
    bg, du = rng.uniform(-5, 80, n), rng.uniform(0, 40, n)
    a, b = m.get_segment_list(pl, lv, bg, du), one.get_segment_list(pl, lv, bg, du)
    assert np.array_equal(a.offsets, b.offsets) and np.array_equal(a.sn, b.sn) and np.array_equal(a.time, b.time)
    m.close()
    one.close()

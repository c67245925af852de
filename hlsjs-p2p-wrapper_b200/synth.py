"""Synthetic playlists and query batches for the BASELINE.json configs.

One integer-only 32-bit generator (mulberry32: imul, shifts, xors) drives everything, so
the same streams can be regenerated in JavaScript (`Math.imul`, `>>> 0`) and C.  All
doubles are dyadic rationals (ticks of 1/1024 s), so every `start` column -- a running sum
of durations -- is exact in fp64 whatever the summation order.  Reference data shape:
`hls.levels[l].details.fragments[i] = {sn, start, duration}` (media-map.js:41-48).
"""
import numpy as np

from .tables import FragmentTables, QueryBatch, TRACK_PARSED

TICK = 1024  # ticks per second
_M32 = np.uint64(0xFFFFFFFF)


def _imul(a, b):
    return (a * b) & _M32


class Mulberry32:
    """Vectorised mulberry32; `u32(n)` returns the next n outputs of the JS generator."""

    def __init__(self, seed):
        self.a = np.uint64(seed & 0xFFFFFFFF)

    def u32(self, n):
        # uint32 arrays: numpy's multiply and add wrap modulo 2^32, which is Math.imul / >>> 0
        k = np.arange(1, n + 1, dtype=np.uint32)
        a = k * np.uint32(0x6D2B79F5) + np.uint32(int(self.a))
        self.a = np.uint64((int(self.a) + n * 0x6D2B79F5) & 0xFFFFFFFF)
        t = (a ^ (a >> np.uint32(15))) * (a | np.uint32(1))
        t = (t + (t ^ (t >> np.uint32(7))) * (t | np.uint32(61))) ^ t
        return (t ^ (t >> np.uint32(14))).astype(np.uint64)          # uint64 holding u32 values

    def below(self, n, bound):
        """n integers uniform in [0, bound) (modulo bias is irrelevant here)."""
        return self.u32(n) % np.uint64(bound)


def _stream(seed, k):
    return Mulberry32((seed * 0x9E3779B1 + k * 0x85EBCA6B + 0x1234567) & 0xFFFFFFFF)


def _vod_tracks(rng, lens):
    """Durations uniform in {2 .. 10 s} in 1/1024 s steps; start = running sum from 0."""
    lens = np.asarray(lens, np.int64)
    off = np.zeros(len(lens) + 1, np.int64)
    np.cumsum(lens, out=off[1:])
    nf = int(off[-1])
    dur_t = (2 * TICK + rng.below(nf, 8 * TICK + 1)).astype(np.int64)   # ticks
    csum = np.cumsum(dur_t)
    excl = csum - dur_t                                      # running sum before fragment i
    first = np.minimum(off[:-1], max(nf - 1, 0))
    track_base = excl[first] if nf else np.zeros(len(lens), np.int64)
    start_t = excl - np.repeat(track_base, lens)             # restart at 0 per track
    total_t = np.zeros(len(lens), np.int64)
    nz = lens > 0
    total_t[nz] = csum[off[1:][nz] - 1] - track_base[nz]
    sn = (np.arange(nf, dtype=np.int64) - np.repeat(off[:-1], lens)).astype(np.int32)
    return off, start_t.astype(np.float64) / TICK, sn, dur_t.astype(np.float64) / TICK, total_t


def _adversarial(rng, n, tables, track, begin, dur):
    """Edge queries: exact boundary hits, rounding sums, negatives, NaN, Inf, -0."""
    kind = rng.below(n, 10).astype(np.int64)
    pick = rng.u32(n).astype(np.int64)
    off = tables.track_off.astype(np.int64)
    o, ln = off[track], off[track + 1] - off[track]
    has = ln > 0
    s = np.zeros(n)
    if has.any() and len(tables.start):
        s[has] = tables.start[o[has] + pick[has] % ln[has]]
    nan, inf = float("nan"), float("inf")
    with np.errstate(invalid="ignore"):
        # 0: begin == a start exactly; 1: begin + duration == a start exactly; 2: a sum that rounds
        # (0.1 + 0.2 style operands are not dyadic); 3: negative duration -> empty; 4, 5: NaN;
        # 6: +Inf; 7: -Inf + Inf = NaN upper bound; 8: -0; 9: -Inf with a huge duration
        b = np.select([kind == 0, kind == 1, kind == 2, kind == 3, kind == 4, kind == 5, kind == 6,
                       kind == 7, kind == 8],
                      [s, s - 30.0, s - 0.3, s, np.full(n, nan), s, s, np.full(n, -inf), np.full(n, -0.0)],
                      default=-inf)
        d = np.select([kind == 0, kind == 1, kind == 2, kind == 3, kind == 4, kind == 5, kind == 6,
                       kind == 7, kind == 8],
                      [np.full(n, 30.0), np.full(n, 30.0), np.full(n, 0.1 + 0.2), np.full(n, -1.0),
                       np.full(n, 30.0), np.full(n, nan), np.full(n, inf), np.full(n, inf), s],
                      default=1e308)
    begin[:] = b
    dur[:] = d


def _queries(rng, tables, playlist, level, total_t, adversarial=True):
    """Query mix of SURVEY 8(d) config 2 for already chosen (playlist, level) pairs."""
    n = len(level)
    track = tables.playlist_base[playlist].astype(np.int64) + level
    tt = total_t[track].astype(np.uint64)
    span = tt + tt // np.uint64(10)                                  # 110 % of the timeline
    begin_t = ((rng.u32(n) * span) >> np.uint64(32)).astype(np.int64) - (tt // np.uint64(20)).astype(np.int64)
    begin = begin_t.astype(np.float64) / TICK
    mix = rng.below(n, 100).astype(np.int64)
    r = rng.u32(n)
    look = (30 * TICK - TICK + (r % np.uint64(2 * TICK + 1)).astype(np.int64)).astype(np.float64) / TICK
    wide = (r % np.uint64(300 * TICK + 1)).astype(np.float64) / TICK
    dur = np.where(mix < 70, look, np.where(mix < 90, wide, 0.0))
    if adversarial:
        idx = np.nonzero(mix >= 95)[0]
        if len(idx):
            b, d = begin[idx].copy(), dur[idx].copy()
            _adversarial(rng, len(idx), tables, track[idx], b, d)
            begin[idx], dur[idx] = b, d
    return QueryBatch(playlist.astype(np.int32), level.astype(np.int32), begin, dur)


def config2(seed=2, n_levels=3, n_frag=1000, n_queries=1_000_000, adversarial=True):
    """VOD playlist, 3 levels x 1k fragments, 10^6 random queries (BASELINE.json configs[1])."""
    rng = _stream(seed, 0)
    off, start, sn, dur, total_t = _vod_tracks(rng, [n_frag] * n_levels)
    tables = FragmentTables(
        np.asarray([0, n_levels], np.int32), off, np.full(n_levels, TRACK_PARSED, np.uint8),
        np.full(n_levels, 2, np.int32), start, sn, dur).validate()
    q = _stream(seed, 1)
    level = q.below(n_queries, n_levels).astype(np.int64)
    playlist = np.zeros(n_queries, np.int64)
    return tables, _queries(q, tables, playlist, level, total_t, adversarial)


def config3(seed=3, n_players=100_000, n_levels=3, window=6):
    """Live sliding windows: sn offset uniform in [0, 2^31 - 7), 6-fragment window.  Per
    player one query over the whole window and one over the next 30 s."""
    rng = _stream(seed, 0)
    nt = n_players * n_levels
    nf = nt * window
    target = np.asarray([2, 4, 6, 10], np.int64)[rng.below(n_players, 4).astype(np.int64)] * TICK
    sn0 = rng.below(n_players, (1 << 31) - 7).astype(np.int64)
    t0 = rng.below(n_players, 86400 * TICK).astype(np.int64)
    jitter = rng.below(nf, TICK // 2 + 1).astype(np.int64) - TICK // 4
    i = np.tile(np.arange(window, dtype=np.int64), nt)
    per_track_target = np.repeat(target, n_levels * window)
    start_t = np.repeat(t0, n_levels * window) + i * per_track_target + jitter
    sn = (np.repeat(sn0, n_levels * window) + i).astype(np.int32)
    tables = FragmentTables(
        (np.arange(n_players + 1, dtype=np.int64) * n_levels).astype(np.int32),
        np.arange(nt + 1, dtype=np.int64) * window, np.full(nt, TRACK_PARSED, np.uint8),
        np.full(nt, 1, np.int32), start_t.astype(np.float64) / TICK, sn,
        per_track_target.astype(np.float64) / TICK).validate()
    q = _stream(seed, 1)
    playlist = np.repeat(np.arange(n_players, dtype=np.int64), 2)
    level = np.repeat(q.below(n_players, n_levels).astype(np.int64), 2)
    first = tables.start[(playlist * n_levels + level) * window]
    begin = first.copy()
    dur = np.empty(2 * n_players)
    dur[0::2] = (target * window).astype(np.float64) / TICK + 1.0     # the whole window
    dur[1::2] = 30.0                                                  # the next 30 s
    begin[1::2] += (q.below(n_players, 20 * TICK).astype(np.float64) / TICK)
    return tables, QueryBatch(playlist.astype(np.int32), level.astype(np.int32), begin, dur)


def config4(seed=4, n_levels=8, n_frag=10_000, n_views=1_000_000, perturbed_pct=1):
    """8 levels x 10k fragments; views drawn as getSegmentList hits (uniform level, index);
    returns (tables, views dict, perturbed copy) for pack -> parse -> isEqual."""
    rng = _stream(seed, 0)
    off, start, sn, dur, _ = _vod_tracks(rng, [n_frag] * n_levels)
    sn = sn + np.int32(1000)
    tables = FragmentTables(
        np.asarray([0, n_levels], np.int32), off, np.full(n_levels, TRACK_PARSED, np.uint8),
        np.full(n_levels, 2, np.int32), start, sn, dur).validate()
    q = _stream(seed, 1)
    level = q.below(n_views, n_levels).astype(np.int64)
    idx = q.below(n_views, n_frag).astype(np.int64)
    url = q.below(n_views, 2).astype(np.int32)
    views = {"level": level.astype(np.int32), "urlId": url, "sn": sn[off[level] + idx],
             "time": start[off[level] + idx]}
    other = {k: v.copy() for k, v in views.items()}
    which = q.below(n_views, 100).astype(np.int64) < perturbed_pct
    field = q.below(n_views, 3).astype(np.int64)
    other["level"][which & (field == 0)] += 1
    other["urlId"][which & (field == 1)] ^= 1
    other["sn"][which & (field == 2)] += 1535
    return tables, views, other


def config5(seed=5, n_playlists=1_000_000, n_levels=3, queries_per_playlist=100,
            adversarial=True):
    """Many playlists, fragments per level piecewise log-uniform in [6, 1024], queries
    grouped by playlist (BASELINE.json configs[4]; scale n_playlists down for tests)."""
    rng = _stream(seed, 0)
    nt = n_playlists * n_levels
    octave = rng.below(nt, 8).astype(np.int64)
    base = 6 << octave
    lens = np.minimum(base + (rng.u32(nt).astype(np.int64) % base), 1024)
    off, start, sn, dur, total_t = _vod_tracks(rng, lens)
    tables = FragmentTables(
        (np.arange(n_playlists + 1, dtype=np.int64) * n_levels).astype(np.int32), off,
        np.full(nt, TRACK_PARSED, np.uint8), np.full(nt, 2, np.int32), start, sn, dur).validate()
    q = _stream(seed, 1)
    nq = n_playlists * queries_per_playlist
    playlist = np.repeat(np.arange(n_playlists, dtype=np.int64), queries_per_playlist)
    level = q.below(nq, n_levels).astype(np.int64)
    return tables, _queries(q, tables, playlist, level, total_t, adversarial)


# BASELINE.json configs[4] at full size, as every script, test and bench leg builds it: 20 chunks of
# 50 000 playlists, chunk c drawn from seed CONFIG5_SEED0 + c (a chunk is generated on its own, so a
# shard needs only the chunks its playlist range touches).  Global playlist id = c * 50 000 + local id.
CONFIG5_CHUNK = 50_000
CONFIG5_CHUNKS = 20
CONFIG5_SEED0 = 1000
CONFIG5_QUERIES_PER_PLAYLIST = 100


def config5_chunk(c, n_playlists=CONFIG5_CHUNK):
    """Chunk c of the full-size configs[4] corpus: (tables, queries) with chunk-local playlist ids."""
    return config5(seed=CONFIG5_SEED0 + c, n_playlists=n_playlists,
                   queries_per_playlist=CONFIG5_QUERIES_PER_PLAYLIST)


def config5_chunk_fragments_per_playlist(c, n_playlists=CONFIG5_CHUNK, n_levels=3):
    """Fragments per playlist of chunk c without building its tables (the first two draws of
    config5's generator): what shard.plan needs to place the shard boundaries."""
    rng = _stream(CONFIG5_SEED0 + c, 0)
    nt = n_playlists * n_levels
    octave = rng.below(nt, 8).astype(np.int64)
    base = 6 << octave
    lens = np.minimum(base + (rng.u32(nt).astype(np.int64) % base), 1024)
    return lens.reshape(n_playlists, n_levels).sum(axis=1)


def hls_mock_tables(level_number=3, defined_level=1, live=False):
    """SoA form of the reference fixture HlsMock(levelNumber, live, definedLevel)
    (test/mocks/hls.js:5-45): sn 25..199, start = 10*sn, no duration, 2 URLs per level."""
    levels = []
    frags = [{"sn": f, "start": f * 10} for f in range(25, 200)]
    for i in range(level_number):
        lvl = {"url": ["http://foo.bar/%d/0/playlist.m3u8" % i, "http://foo.bar/%d/1/playlist.m3u8" % i]}
        if live is not None and i == defined_level:
            lvl["details"] = {"live": live, "fragments": frags}
        levels.append(lvl)
    return FragmentTables.from_hls_levels([levels if level_number > 0 else None])


# ---- media playlists in text form (input shape per a13) -------------------
_NASTY = [
    b"#EXTM3U\n#EXT-X-MEDIA-SEQUENCE:7\n#EXTINF:10,title #EXTINF:4,\r\nseg.ts\n#EXTINF:.5\n"
    b"#EXT-X-KEY:METHOD=NONE#EXTINF:7,\n#EXTINF:.,\nx.ts\n#EXTINF:1.2.3,\n#EXTINF:abc,\n#EXT-X-ENDLIST",
    b"#EXTINF:3#EXT-X-ENDLIST\n#EXT-X-ENDLISTING#EXTINF:2,\n#EXT-X-MEDIA-SEQUENCE:12#EXTINF:4,\n"
    b"#EXT-X-MEDIA-SEQUENCE:x\n#EXTINF:5.,\n",
    b"#EXT-X-PROGRAM-DATE-TIME:2016-01-01T00:00:00Z\xe2\x80\xa8#EXTINF:6,\n"
    b"#EXT-X-KEY:URI=\"a#EXT-X-MEDIA-SEQUENCE:99\"\n#EXTINF:0.000,\nurl#EXTINF:8,\n#EXTINF:9.98458,|#EXTINF:3\n",
    b"#EXT-X-MEDIA-SEQUENCE:2147483640\n#EXTINF:00010.500000000000000000000,\n#EXTINF:0.1234567890123456789,\n"
    b"#EXTINF:123456789012345678,\n#EXTINF:0.30000000000000004,\n#EXTINF:9007199254740993,\n",
    b"",
    b"no tags here at all\n",
    b"#EXT-X-MEDIA-SEQUENCE:5\n#EXTINF:2,\na\n#EXTINF:2,\nb\n#EXT-X-MEDIA-SEQUENCE:100\n#EXTINF:2,\nc\n"
    b"#EXT-X-DISCONTINUITY\n#EXT-X-BYTERANGE:100@0\n#EXTINF:2,\nd\n#EXT-X-MEDIA-SEQUENCE:3\n",
    b"#EXTINF:4,\n" * 3 + b"#EXT-X-MEDIA-SEQUENCE:",
    b"#EXTINF:",
    b"#" * 40 + b"#EXTINF:1," + b"#" * 9,
]


def m3u8_corpus(seed=8, n_playlists=200, n_levels=3, frags=(3, 12), nasty=True, unparsed_every=0):
    """Level playlists as text -> (playlist_base int32[P+1], track_state uint8[T], texts list[bytes]).
    Durations are printed in the formats segmenters use (integers, 3/5/6 decimals); every few
    tracks use CRLF, drop the final newline, or carry no #EXT-X-MEDIA-SEQUENCE.  With `nasty`, hand
    written tracks exercising the parser's corners are mixed in."""
    rng = np.random.default_rng(seed)
    nt = n_playlists * n_levels
    lens = rng.integers(frags[0], frags[1] + 1, nt)
    texts, state = [], np.full(nt, TRACK_PARSED, np.uint8)
    fmts = ("%d", "%.3f", "%.5f", "%.6f")
    for t in range(nt):
        if unparsed_every and t % unparsed_every == unparsed_every - 1:
            state[t] = t % 2                          # missing or unparsed: text must be ignored
        if nasty and t % 11 == 5:
            texts.append(_NASTY[(t // 11) % len(_NASTY)])
            continue
        nl = "\r\n" if t % 7 == 3 else "\n"
        fmt = fmts[t % 4]
        sn0 = int(rng.integers(0, 2_000_000_000))
        d = rng.integers(2000, 10001, int(lens[t])) / 1000.0
        head = ["#EXTM3U", "#EXT-X-VERSION:3", "#EXT-X-TARGETDURATION:10"]
        if t % 13 != 12:
            head.append("#EXT-X-MEDIA-SEQUENCE:%d" % sn0)
        body = []
        for i, x in enumerate(d):
            body.append("#EXTINF:" + (fmt % (round(x) if fmt == "%d" else x)) + ",")
            body.append("fileSequence%d.ts" % (sn0 + i))
        if t % 5 == 0:
            body.append("#EXT-X-ENDLIST")
        s = nl.join(head + body) + ("" if t % 9 == 4 else nl)
        texts.append(s.encode())
    base = (np.arange(n_playlists + 1, dtype=np.int64) * n_levels).astype(np.int32)
    return base, state, texts


def m3u8_block(seed=9, n_tracks=1000, n_frag=800):
    """A large, regular corpus built quickly (numpy): n_tracks VOD playlists of n_frag fragments with
    5-decimal durations.  -> (text uint8[N], text_off int64[T+1])"""
    rng = np.random.default_rng(seed)
    nf = n_tracks * n_frag
    ms = rng.integers(200000, 1000001, nf)            # 2.00000 .. 10.00000 s in 1e-5 steps
    line = np.frombuffer(b"#EXTINF:0.00000,\nseg0000000.ts\n", np.uint8)
    rows = np.tile(line, (nf, 1))
    for k, col in enumerate((8, 10, 11, 12, 13, 14)):  # digits of d.ddddd
        div = 10 ** (5 - k)
        rows[:, col] = 48 + (ms // div) % 10
    idx = np.arange(nf) % n_frag
    for k in range(7):
        rows[:, 20 + k] = 48 + (idx // 10 ** (6 - k)) % 10
    head = b"#EXTM3U\n#EXT-X-TARGETDURATION:10\n#EXT-X-MEDIA-SEQUENCE:1000\n"
    tail = b"#EXT-X-ENDLIST\n"
    per = len(head) + n_frag * len(line) + len(tail)
    out = np.empty((n_tracks, per), np.uint8)
    out[:, :len(head)] = np.frombuffer(head, np.uint8)
    out[:, len(head):len(head) + n_frag * len(line)] = rows.reshape(n_tracks, -1)
    out[:, per - len(tail):] = np.frombuffer(tail, np.uint8)
    return out.reshape(-1), np.arange(n_tracks + 1, dtype=np.int64) * per


def sample_doubles(n, seed):
    """Finite doubles of every kind Number::toString lays out differently (any exponent, playlist
    times, powers of two, subnormals, the classic hard cases), some negated."""
    rng = np.random.default_rng(seed)
    parts = [rng.integers(0, 2 ** 63, n // 4, dtype=np.int64).view(np.float64),          # any exponent
             rng.integers(0, 10 ** 8, n // 4) / 1000.0,                                    # playlist times
             rng.integers(0, 86400 * 1024, n // 4) / 1024.0 + rng.integers(0, 1000, n // 4) * 1e-7,
             np.ldexp(1.0, rng.integers(-1074, 1024, n // 8)),                             # powers of two
             rng.integers(0, 2 ** 52, n // 8, dtype=np.int64).view(np.float64),            # subnormals
             np.asarray([0.1 + 0.2, 1e21, 1e-7, 123e-20, 2.0 ** -24, 5e-324, 1.7976931348623157e308, 1e20,
                         1e-6, 999999999999999900000.0, 9007199254740993.0, 4.35, 0.000001, 1.5e300])]
    x = np.concatenate(parts)
    x = x[np.isfinite(x)]
    return np.concatenate([x, -x[:1000]])


def m3u8_token_soup(rng):
    """Text glued from the tokens the expression cares about, with no regard for the HLS grammar."""
    tokens = ["#EXTINF:", "#EXT-X-MEDIA-SEQUENCE:", "#EXT-X-ENDLIST", "#EXT-X-KEY:", "#EXT-X-PROGRAM-DATE-TIME:",
              "#EXT-X-TARGETDURATION:", "#EXT-X-DISCONTINUITY", "#EXT-X-BYTERANGE:", "#EXTM3U", "#EXT", "#", "\n", "\n",
              "\r\n", "\r", ",", ".", "|", "@", "seg.ts", "9.98458", "10", "0.5", ".5", "5.", "1.2.3", "007", " ", "x",
              " ", "12345678901234567", "0.1234567", "2147483640", "#EXT-X-", ":"]
    return "".join(tokens[i] for i in rng.integers(0, len(tokens), int(rng.integers(1, 60))))

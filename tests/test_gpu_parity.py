"""Bit-exact parity of the CUDA path (through the C ABI) against the CPU oracle on seeded
synthetic inputs: counts, offsets, per-hit sn (int) and time (fp64 bit pattern) in order,
statuses; both kernels (search / filter); edge cases; the BASELINE.json configs at sizes
the oracle finishes in seconds, plus size-independent properties at full size."""
import ctypes

import numpy as np
import pytest

import oracle
import hlsjs_p2p_wrapper_b200 as pk
from hlsjs_p2p_wrapper_b200 import synth, _native as N
from hlsjs_p2p_wrapper_b200.tables import FragmentTables, QueryBatch, TRACK_PARSED

pytestmark = pytest.mark.gpu


def oracle_list(t, q, threads=8):
    return oracle.get_segment_list_batch(t.playlist_base, t.track_off, t.track_state, t.start, t.sn,
                                         q.playlist, q.level, q.begin, q.duration, threads=threads)


def assert_same(res, ref, what=""):
    off, status, sn, tm = ref
    assert res.total_hits == off[-1], what
    assert np.array_equal(res.offsets.astype(np.int64), off), what
    if res.status is not None:
        assert np.array_equal(res.status, status), what
    assert np.array_equal(res.sn, sn), what
    assert np.array_equal(res.time.view(np.uint64), tm.view(np.uint64)), what   # bit patterns
    assert res.n_unparsed == int((status == 1).sum()) and res.n_missing == int((status == 2).sum()), what


def run_both_paths(ctx, t, q, paths=(N.PATH_SEARCH, N.PATH_FILTER)):
    m = ctx.upload(t)
    ref = oracle_list(t, q)
    for path in paths:
        m.set_path(path)
        res = m.get_segment_list(q.playlist, q.level, q.begin, q.duration)
        assert res.path == path and res.kernel_launches >= 1
        assert_same(res, ref, "path %d" % path)
    m.close()
    return ref


# ---- config 2: VOD, 3 levels x 1k fragments ------------------------------------------------
@pytest.mark.parametrize("nq", [1, 3, 4, 31, 255, 1023, 1024, 1025, 4099, 100_000])
def test_config2_sizes(ctx, nq):
    t, q = synth.config2(seed=20 + nq % 7, n_queries=nq)
    run_both_paths(ctx, t, q)


def test_config2_full_size_search(ctx):
    """BASELINE.json configs[1] at full size (10^6 queries), bit-exact against the oracle."""
    t, q = synth.config2()
    m = ctx.upload(t)
    info = m.info()
    assert info["all_monotone"] == 1 and info["stage_all"] == 1
    res = m.get_segment_list(q.playlist, q.level, q.begin, q.duration)
    assert res.path == N.PATH_SEARCH and res.kernel_launches >= 1
    assert_same(res, oracle_list(t, q))
    # range form agrees with the list form
    first, count, status = m.get_segment_range(q.playlist, q.level, q.begin, q.duration)
    assert np.array_equal(count, np.diff(res.offsets))
    rf, rc, rs = oracle.get_segment_range_batch(t.playlist_base, t.track_off, t.track_state, t.start,
                                                q.playlist, q.level, q.begin, q.duration)
    assert np.array_equal(first, rf) and np.array_equal(count, rc) and np.array_equal(status, rs)
    m.close()


def test_empty_batch_and_empty_map(ctx):
    t, q = synth.config2(n_queries=16)
    m = ctx.upload(t)
    res = m.get_segment_list(q.playlist[:0], q.level[:0], q.begin[:0], q.duration[:0])
    assert res.total_hits == 0 and list(res.offsets) == [0] and len(res.sn) == 0
    m.close()
    empty = FragmentTables.from_hls_levels([None, []])          # levels undefined / []
    m = ctx.upload(empty)
    res = m.get_segment_list([0, 1, 2, -1], [0, 0, 0, 0], [0.0] * 4, [1.0] * 4)
    assert list(res.status) == [2, 2, 2, 2] and res.n_missing == 4 and res.total_hits == 0
    m.close()


def test_ragged_levels_and_states(ctx):
    """Playlists with 0..4 levels, missing / unparsed / empty / one-fragment tracks."""
    lv = lambda fr: {"url": ["a", "b"], "details": {"fragments": fr}}
    f = lambda n, s0=0.0: [{"sn": 100 + i, "start": s0 + 2.5 * i, "duration": 2.5} for i in range(n)]
    playlists = [
        [lv(f(5)), {"url": ["a"]}, lv([]), lv(f(1, 7.0))],
        None,
        [],
        [None, lv(f(40, -30.0))],
        [lv(f(3))],
    ]
    t = FragmentTables.from_hls_levels(playlists)
    rng = np.random.default_rng(5)
    nq = 5000
    q = QueryBatch(rng.integers(-1, 6, nq).astype(np.int32), rng.integers(-1, 5, nq).astype(np.int32),
                   rng.integers(-40, 80, nq) / 4.0, rng.integers(-4, 60, nq) / 4.0)
    ref = run_both_paths(ctx, t, q)
    assert set(np.unique(ref[1])) == {0, 1, 2}
    m = ctx.upload(t)
    out, st = m.get_segment_duration(q.playlist, q.level)
    eo, es = oracle.get_segment_duration_batch(t.playlist_base, t.track_off, t.track_state, t.dur0(),
                                               q.playlist, q.level)
    assert np.array_equal(st, es) and np.array_equal(out, eo, equal_nan=True)
    assert set(np.unique(es)) == {0, 1, 2, 3}
    for p in range(len(playlists)):
        l, u = m.get_track_list(p)
        b, e = t.playlist_base[p], t.playlist_base[p + 1]
        el, eu = oracle.get_track_list(t.url_count[b:e])
        assert np.array_equal(l, el) and np.array_equal(u, eu)
    m.close()


def test_special_values_on_monotone_table(ctx):
    """NaN / +-Inf / -0 / duplicates / rounding sums as query operands and table entries."""
    s = 0.1 + 0.2
    start = np.array([-np.inf, -5.0, -0.0, 0.0, 0.0, 0.3, s, np.nextafter(s, 1), 7.0, 7.0, 7.0, 1e308, np.inf, np.inf])
    t = FragmentTables(np.array([0, 1], np.int32), np.array([0, len(start)], np.int64),
                       np.array([TRACK_PARSED], np.uint8), np.array([1], np.int32), start,
                       np.arange(len(start), dtype=np.int32), np.full(len(start), 1.0))
    vals = [np.nan, np.inf, -np.inf, 0.0, -0.0, 0.1, 0.2, 0.3, s, 7.0, -5.0, 1e308, -1e308, 5e-324, 1.7e308]
    b, d = np.meshgrid(vals, vals)
    q = QueryBatch(np.zeros(b.size, np.int32), np.zeros(b.size, np.int32), b.ravel().copy(), d.ravel().copy())
    m = ctx.upload(t)
    assert m.info()["all_monotone"] == 1
    m.close()
    run_both_paths(ctx, t, q)


def test_non_monotone_tables_use_filter(ctx):
    """Unsorted and NaN-bearing start columns: only the filter kernel is legal, AUTO picks it,
    forcing the search path is refused, and results match the reference filter."""
    rng = np.random.default_rng(11)
    lens = [0, 1, 2, 33, 64, 65, 200, 1000]
    cols, sns = [], []
    for i, n in enumerate(lens):
        c = rng.integers(0, 400, n) / 4.0
        if i % 2 == 0 and n > 2:
            c[rng.integers(0, n, max(1, n // 10))] = np.nan
        if i == 6:
            c = np.sort(c)                                      # one sorted track among them
        cols.append(c)
        sns.append(rng.integers(0, 1 << 30, n))
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    t = FragmentTables(np.array([0, 3, 8], np.int32), off, np.full(8, TRACK_PARSED, np.uint8),
                       np.full(8, 1, np.int32), np.concatenate(cols), np.concatenate(sns).astype(np.int32),
                       np.ones(int(off[-1])))
    m = ctx.upload(t)
    info = m.info()
    flags = m.monotone_flags()
    assert info["all_monotone"] == 0
    assert list(flags) == [int(oracle.is_monotone(c)) for c in cols]
    with pytest.raises(N.HlsMapError):
        m.set_path(N.PATH_SEARCH)
    nq = 20000
    q = QueryBatch(rng.integers(0, 2, nq).astype(np.int32), rng.integers(0, 5, nq).astype(np.int32),
                   rng.integers(-20, 420, nq) / 4.0, rng.integers(-8, 200, nq) / 4.0)
    res = m.get_segment_list(q.playlist, q.level, q.begin, q.duration)
    assert res.path == N.PATH_FILTER
    assert_same(res, oracle_list(t, q))
    first, count, status = m.get_segment_range(q.playlist, q.level, q.begin, q.duration)
    assert np.array_equal(count, np.diff(res.offsets))
    assert (status == N.ST_SCATTERED).any()
    m.close()


def test_capacity_contract(ctx):
    t, q = synth.config2(n_queries=5000)
    m = ctx.upload(t)
    off, total = m.count_segments(q.playlist, q.level, q.begin, q.duration)
    ref = oracle_list(t, q)
    assert total == ref[0][-1] and np.array_equal(off.astype(np.int64), ref[0])
    with pytest.raises(N.HlsMapError) as e:
        m.get_segment_list(q.playlist, q.level, q.begin, q.duration, capacity=total - 1)
    assert e.value.code == N.ERR_CAPACITY
    res = m.get_segment_list(q.playlist, q.level, q.begin, q.duration, capacity=total)
    assert_same(res, ref)
    m.close()


def test_repeated_launches_share_scratch(ctx):
    """Back-to-back batches of different sizes through one context (epoch-tagged scratch)."""
    t, q = synth.config2(n_queries=30000)
    m = ctx.upload(t)
    ref = oracle_list(t, q)
    for n in (30000, 7, 2048, 30000, 1, 29999):
        sub = QueryBatch(q.playlist[:n], q.level[:n], q.begin[:n], q.duration[:n])
        res = m.get_segment_list(sub.playlist, sub.level, sub.begin, sub.duration)
        assert np.array_equal(res.offsets.astype(np.int64), ref[0][:n + 1])
        assert np.array_equal(res.sn, ref[2][:ref[0][n]])
    m.close()


# ---- config 3: live windows ------------------------------------------------------------------
def test_config3_live_windows(ctx):
    t, q = synth.config3(n_players=20_000)
    ref = run_both_paths(ctx, t, q)
    counts = np.diff(ref[0])
    assert (counts[0::2] == 6).all()            # the whole-window query keeps all 6 fragments
    m = ctx.upload(t)
    assert m.info()["stage_all"] == 0
    out, st = m.get_segment_duration(q.playlist, q.level)
    eo, es = oracle.get_segment_duration_batch(t.playlist_base, t.track_off, t.track_state, t.dur0(),
                                               q.playlist, q.level)
    assert np.array_equal(out, eo) and np.array_equal(st, es)
    m.close()


# ---- config 5: many playlists, queries grouped by playlist -------------------------------------
def test_config5_scaled(ctx):
    t, q = synth.config5(n_playlists=3000, queries_per_playlist=100)
    run_both_paths(ctx, t, q, paths=(N.PATH_SEARCH,))
    t, q = synth.config5(seed=6, n_playlists=200, queries_per_playlist=50)
    run_both_paths(ctx, t, q)


def test_config5_shuffled_queries_span_beyond_window(ctx):
    """Un-grouped queries: a tile's tracks are far apart, so the search falls back to
    probing L2/HBM directly instead of a staged window."""
    t, q = synth.config5(seed=7, n_playlists=4000, queries_per_playlist=10)
    perm = np.random.default_rng(3).permutation(len(q))
    q = QueryBatch(q.playlist[perm], q.level[perm], q.begin[perm], q.duration[perm])
    run_both_paths(ctx, t, q, paths=(N.PATH_SEARCH,))


@pytest.mark.parametrize("window_kb", [2, 16, 24])
def test_config5_tracks_spanning_several_windows(ctx, window_kb, monkeypatch):
    """A staging window (HLSMAP_STAGE_WINDOW_KB, read at upload) smaller than a tile's tracks:
    the tile is searched in several staging passes, tracks longer than the window and tiles
    spread over more than a few windows probe global memory -- same output in every case."""
    monkeypatch.setenv("HLSMAP_STAGE_WINDOW_KB", str(window_kb))
    t, q = synth.config5(seed=11, n_playlists=300, queries_per_playlist=100)
    m = ctx.upload(t)
    assert m.info()["stage_all"] == 0
    ref = oracle_list(t, q)
    assert_same(m.get_segment_list(q.playlist, q.level, q.begin, q.duration), ref, "grouped")
    m.close()
    # half-shuffled: runs of grouped queries interleaved with far-away ones
    rng = np.random.default_rng(window_kb)
    perm = np.arange(len(q))
    sel = rng.random(len(q)) < 0.3
    perm[sel] = rng.permutation(perm[sel])
    q2 = QueryBatch(q.playlist[perm], q.level[perm], q.begin[perm], q.duration[perm])
    m = ctx.upload(t)
    assert_same(m.get_segment_list(q2.playlist, q2.level, q2.begin, q2.duration), oracle_list(t, q2), "mixed")
    m.close()


def test_huge_counts_take_the_cooperative_expansion(ctx):
    """Queries that keep whole 10k-fragment tracks (warp output run > 2048 entries)."""
    t, _, _ = synth.config4(n_views=8)
    nq = 3000
    rng = np.random.default_rng(2)
    q = QueryBatch(np.zeros(nq, np.int32), rng.integers(0, 8, nq).astype(np.int32),
                   rng.integers(-100, 70000, nq).astype(np.float64),
                   np.where(rng.random(nq) < 0.1, 1e6, 30.0))
    run_both_paths(ctx, t, q)


# ---- device-pointer entry point, properties at full size ---------------------------------------
def test_device_pointer_variant_full_size_properties(ctx):
    import torch
    t, q = synth.config2()
    m = ctx.upload(t)
    dev = torch.device("cuda:0")
    d_lv = torch.from_numpy(q.level).to(dev)
    d_bg = torch.from_numpy(q.begin).to(dev)
    d_du = torch.from_numpy(q.duration).to(dev)
    nq = len(q)
    cap = 20 * nq
    d_off = torch.empty(nq + 1, dtype=torch.int32, device=dev)
    d_sn = torch.empty(cap, dtype=torch.int32, device=dev)
    d_tm = torch.empty(cap, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    s = torch.cuda.current_stream().cuda_stream
    res = m.get_segment_list_dev(nq, 0, d_lv.data_ptr(), d_bg.data_ptr(), d_du.data_ptr(),
                                 d_off.data_ptr(), d_sn.data_ptr(), d_tm.data_ptr(), cap,
                                 stream=s, sync=True)
    off = d_off.cpu().numpy().astype(np.int64)
    H = int(res.total_hits)
    assert off[0] == 0 and off[-1] == H and (np.diff(off) >= 0).all()
    sn, tm = d_sn[:H].cpu().numpy(), d_tm[:H].cpu().numpy()
    # every hit satisfies the predicate with its own query, and runs are sorted by time
    qi = np.repeat(np.arange(nq), np.diff(off))
    with np.errstate(invalid="ignore"):
        ub = q.begin + q.duration
    assert (q.begin[qi] <= tm).all() and (tm <= ub[qi]).all()
    lvl = q.level[qi].astype(np.int64)
    assert np.array_equal(t.start[t.track_off[lvl] + sn], tm)      # sn indexes the same fragment
    # asynchronous form + last_result
    m.get_segment_list_dev(nq, 0, d_lv.data_ptr(), d_bg.data_ptr(), d_du.data_ptr(),
                           d_off.data_ptr(), d_sn.data_ptr(), d_tm.data_ptr(), cap, stream=s)
    again = m.last_result(s)
    assert again.total_hits == H
    assert_same_dev = oracle_list(t, q)
    assert np.array_equal(off, assert_same_dev[0]) and np.array_equal(sn, assert_same_dev[2])
    assert np.array_equal(tm.view(np.uint64), assert_same_dev[3].view(np.uint64))
    m.close()


# ---- host-pointer pipeline (chunked H2D / kernel / D2H) ------------------------------------------
def test_host_pipeline_many_hits_falls_back_and_stays_exact(ctx):
    """> 16 hits per query on average: the pipeline's device-side guess overflows, the call falls
    back to the exact-count path, and the result is still bit-exact."""
    t, _ = synth.config2(n_queries=4)
    nq = 300_000
    rng = np.random.default_rng(8)
    q = QueryBatch(np.zeros(nq, np.int32), rng.integers(0, 3, nq).astype(np.int32),
                   rng.integers(0, 5000, nq).astype(np.float64), np.full(nq, 200.0))
    m = ctx.upload(t)
    ref = oracle_list(t, q)
    assert ref[0][-1] > 16 * nq + 1024
    res = m.get_segment_list(q.playlist, q.level, q.begin, q.duration, capacity=int(ref[0][-1]))
    assert_same(res, ref)
    m.close()


def test_host_pipeline_capacity_and_statuses(ctx):
    """Chunked path: caller capacity one short -> CAPACITY error with the needed total; statuses
    and unparsed / missing tallies accumulate across chunks."""
    t = synth.hls_mock_tables(3, 1, False)              # level 1 parsed, 0 and 2 unparsed
    nq = 600_000
    rng = np.random.default_rng(9)
    q = QueryBatch(np.zeros(nq, np.int32), rng.integers(-1, 5, nq).astype(np.int32),
                   rng.integers(0, 2500, nq).astype(np.float64), rng.integers(0, 60, nq).astype(np.float64))
    m = ctx.upload(t)
    ref = oracle_list(t, q)
    total = int(ref[0][-1])
    res = m.get_segment_list(q.playlist, q.level, q.begin, q.duration, capacity=total)
    assert res.kernel_launches >= 2                      # it really went through the chunks
    assert_same(res, ref)
    with pytest.raises(N.HlsMapError) as e:
        m.get_segment_list(q.playlist, q.level, q.begin, q.duration, capacity=total - 1)
    assert e.value.code == N.ERR_CAPACITY and str(total) in e.value.message
    m.close()


def test_overlapping_batches_on_disjoint_buffers(ctx):
    """hlsmap_ctx_set_batch_overlap: consecutive batches with their own buffers may overlap on
    the GPU (the next starts while the previous drains); every batch must still be bit-exact,
    and the library's double-buffered scratch must survive many launches in a row."""
    import torch
    dev = torch.device("cuda:0")
    t, _ = synth.config2(n_queries=4)
    m = ctx.upload(t)
    stream = torch.cuda.Stream(device=dev)
    sets = []
    # (the 10^6-query batches take the 1024-query-tile variant the library picks for large
    #  overlapped batches, the others 512-query tiles: both kinds alternate on the same scratch)
    for i, nq in enumerate((200_000, 1_000_003, 70_001, 333_333, 1_000, 1_000_000, 150_000)):
        _, q = synth.config2(seed=40 + i, n_queries=nq)
        ref = oracle_list(t, q)
        H = int(ref[0][-1])
        sets.append({"q": q, "ref": ref, "H": H,
                     "pl": torch.from_numpy(q.playlist).to(dev), "lv": torch.from_numpy(q.level).to(dev),
                     "bg": torch.from_numpy(q.begin).to(dev), "du": torch.from_numpy(q.duration).to(dev),
                     "off": torch.empty(nq + 1, dtype=torch.int32, device=dev),
                     "sn": torch.empty(max(H, 1), dtype=torch.int32, device=dev),
                     "tm": torch.empty(max(H, 1), dtype=torch.float64, device=dev)})
    torch.cuda.synchronize()
    ctx.set_batch_overlap(True)
    try:
        for rep in range(4):                       # 28 launches back to back, no host sync between
            for s in sets:
                with torch.cuda.stream(stream):     # on the batches' own stream: ordered before the launch
                    s["off"].fill_(-1)
                m.get_segment_list_dev(len(s["q"]), s["pl"].data_ptr(), s["lv"].data_ptr(),
                                       s["bg"].data_ptr(), s["du"].data_ptr(), s["off"].data_ptr(),
                                       s["sn"].data_ptr(), s["tm"].data_ptr(), s["H"],
                                       stream=stream.cuda_stream)
        last = m.last_result(stream.cuda_stream)
        assert last.total_hits == sets[-1]["H"]
        torch.cuda.synchronize()
        for s in sets:
            assert np.array_equal(s["off"].cpu().numpy().astype(np.int64), s["ref"][0])
            assert np.array_equal(s["sn"][:s["H"]].cpu().numpy(), s["ref"][2])
            assert np.array_equal(s["tm"][:s["H"]].cpu().numpy().view(np.uint64), s["ref"][3].view(np.uint64))
    finally:
        ctx.set_batch_overlap(False)
    m.close()


def test_small_batches_behind_big_ones_release_their_scratch_set(ctx):
    """A batch of two tiles launched while a big batch still holds the SMs: the small grid's early
    CTA can finish both tiles before its other CTA is even dispatched.  That late CTA must find the
    scratch set still its own launch's and leave it properly, or the launch after next would wait
    for the set for ever.  A few thousand alternations on two streams, both kernels of the list
    form, every small result compared with the oracle; the watchdog turns a hang into a failure."""
    import threading
    import torch
    dev = torch.device("cuda:0")
    t, _ = synth.config2(n_queries=4)
    m = ctx.upload(t)
    s_small, s_big = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def dev_set(seed, nq):
        _, q = synth.config2(seed=seed, n_queries=nq)
        ref = oracle_list(t, q)
        H = int(ref[0][-1])
        return {"q": q, "ref": ref, "H": H,
                "pl": torch.from_numpy(q.playlist).to(dev), "lv": torch.from_numpy(q.level).to(dev),
                "bg": torch.from_numpy(q.begin).to(dev), "du": torch.from_numpy(q.duration).to(dev),
                "off": torch.empty(nq + 1, dtype=torch.int32, device=dev),
                "sn": torch.empty(max(H, 1), dtype=torch.int32, device=dev),
                "tm": torch.empty(max(H, 1), dtype=torch.float64, device=dev)}

    small = [dev_set(90 + i, nq) for i, nq in enumerate((700, 1024, 513, 1500))]     # 2-3 tiles each
    big = dev_set(99, 400_000)
    # a second context on the same GPU keeps the SMs busy with its own batches meanwhile
    other = pk.Context(0)
    m2 = other.upload(t)

    def launch(mm, s, stream):
        mm.get_segment_list_dev(len(s["q"]), s["pl"].data_ptr(), s["lv"].data_ptr(), s["bg"].data_ptr(),
                                s["du"].data_ptr(), s["off"].data_ptr(), s["sn"].data_ptr(),
                                s["tm"].data_ptr(), s["H"], stream=stream.cuda_stream)

    done = threading.Event()

    def work():
        for path in (N.PATH_SEARCH, N.PATH_FILTER):
            m.set_path(path)
            ctx.set_batch_overlap(path == N.PATH_SEARCH)
            reps = 1500 if path == N.PATH_SEARCH else 40
            for i in range(reps):
                if i % 8 == 0:
                    launch(m2, big, s_big)
                launch(m, small[i % len(small)], s_small)
            torch.cuda.synchronize()
            for s in small:
                assert np.array_equal(s["off"].cpu().numpy().astype(np.int64), s["ref"][0])
                assert np.array_equal(s["sn"][:s["H"]].cpu().numpy(), s["ref"][2])
                assert np.array_equal(s["tm"][:s["H"]].cpu().numpy().view(np.uint64), s["ref"][3].view(np.uint64))
        done.set()

    err = []

    def guarded():
        try:
            work()
        except BaseException as e:      # surfaced below
            err.append(e)
            done.set()

    th = threading.Thread(target=guarded, daemon=True)
    th.start()
    ok = done.wait(timeout=120)
    ctx.set_batch_overlap(False)
    assert ok, "launches hung: a scratch set was never released"
    if err:
        raise err[0]
    m.set_path(N.PATH_AUTO)
    m2.close()
    other.close()
    m.close()


@pytest.mark.parametrize("which", ["small_map", "large_map"])
def test_fused_batches_match_the_oracle_batch_by_batch(ctx, which):
    """Under hlsmap_ctx_set_batch_overlap a batches call resolves runs of up to 32 non-empty batches
    with ONE kernel launch (their tiles numbered consecutively, each batch with its own look-back
    chain, buffers, capacity and status column).  Forty batches of mixed sizes -- more than one launch
    takes, an empty one and a count-only one among them, some with a status column, one with too small a
    capacity -- on a map that fits shared memory and on one that does not: every batch bit-for-bit the
    oracle's answer, identical to what one launch per batch (HLSMAP_NO_FUSE) gives, and a broken batch
    in the middle stops the call exactly there."""
    import torch
    dev = torch.device("cuda:0")
    if which == "small_map":
        t, _ = synth.config2(n_queries=4)
        gen = lambda seed, nq: synth.config2(seed=seed, n_queries=nq)[1]
    else:
        t, q_all = synth.config5(seed=77, n_playlists=3000, queries_per_playlist=40)
        def gen(seed, nq):
            r = np.random.RandomState(seed)
            i = np.sort(r.randint(0, len(q_all), size=nq))      # (grouped by playlist, as the corpus is)
            return QueryBatch(q_all.playlist[i], q_all.level[i], q_all.begin[i], q_all.duration[i])
    m = ctx.upload(t)
    stream = torch.cuda.Stream(device=dev)
    sizes = [30_000, 513, 1, 70_001, 0, 1024, 2_049, 120_000] * 5
    sets = []
    for i, nq in enumerate(sizes):
        q = gen(300 + i, max(nq, 1))
        q = QueryBatch(q.playlist[:nq], q.level[:nq], q.begin[:nq], q.duration[:nq])
        ref = oracle_list(t, q)
        H = int(ref[0][-1])
        count_only = i == 11
        short = i == 19 and H > 8              # room for fewer hits than the batch has
        cap = 0 if count_only else (H // 2 if short else H)
        sets.append({"ref": ref, "H": H, "nq": nq, "cap": cap, "count_only": count_only,
                     "pl": torch.from_numpy(q.playlist).to(dev), "lv": torch.from_numpy(q.level).to(dev),
                     "bg": torch.from_numpy(q.begin).to(dev), "du": torch.from_numpy(q.duration).to(dev),
                     "off": torch.full((nq + 1,), -1, dtype=torch.int32, device=dev),
                     "sn": torch.full((max(H, 1) + 256,), -7, dtype=torch.int32, device=dev),
                     "tm": torch.full((max(H, 1) + 256,), -7.0, dtype=torch.float64, device=dev),
                     "st": torch.full((max(nq, 1),), 99, dtype=torch.uint8, device=dev) if i % 3 == 0 else None})
    def batch(s):
        return {"nq": s["nq"], "d_playlist": s["pl"].data_ptr(), "d_level": s["lv"].data_ptr(),
                "d_begin": s["bg"].data_ptr(), "d_duration": s["du"].data_ptr(), "d_offsets": s["off"].data_ptr(),
                "d_sn_out": 0 if s["count_only"] else s["sn"].data_ptr(),
                "d_time_out": 0 if s["count_only"] else s["tm"].data_ptr(), "capacity": s["cap"],
                "d_status": s["st"].data_ptr() if s["st"] is not None else 0}
    def check():
        torch.cuda.synchronize()
        for s in sets:
            assert np.array_equal(s["off"].cpu().numpy().astype(np.int64), s["ref"][0])
            if s["st"] is not None and s["nq"]:
                assert np.array_equal(s["st"][:s["nq"]].cpu().numpy(), s["ref"][1])
            if s["count_only"]:
                continue
            k = s["cap"]                       # hits beyond the capacity are not written, nothing behind them is touched
            assert np.array_equal(s["sn"][:k].cpu().numpy(), s["ref"][2][:k])
            assert np.array_equal(s["tm"][:k].cpu().numpy().view(np.uint64), s["ref"][3][:k].view(np.uint64))
            assert (s["sn"][k:] == -7).all().item() and (s["tm"][k:] == -7.0).all().item()
    def reset():
        for s in sets:
            s["off"].fill_(-1); s["sn"].fill_(-7); s["tm"].fill_(-7.0)
            if s["st"] is not None:
                s["st"].fill_(99)
        torch.cuda.synchronize()
    ctx.set_batch_overlap(True)
    try:
        launches = {}
        for mode in (1, 1, 2, 1):               # 1: runs of batches fused into one launch; 2: one launch per batch
            ctx.set_batch_overlap(mode)
            reset()
            assert m.get_segment_list_dev_batches([batch(s) for s in sets], stream.cuda_stream) == len(sets)
            last = m.last_result(stream.cuda_stream)
            assert last.total_hits == sets[-1]["H"]
            launches[mode] = last.kernel_launches          # of the whole call
            check()
        assert launches[2] == sum(1 for s in sets if s["nq"] > 0)
        assert launches[1] == 6                             # the six runs between the five empty batches
        ctx.set_batch_overlap(True)
        # a batch that cannot be launched ends the call in front of it, fused or not
        reset()
        bad = dict(batch(sets[5]), d_begin=0)
        arr = (N.Batch * 9)()
        for a, b in zip(arr, [batch(s) for s in sets[:4]] + [bad] + [batch(s) for s in sets[5:9]]):
            a.nq, a.d_playlist, a.d_level, a.d_begin, a.d_duration = b["nq"], b["d_playlist"], b["d_level"], b["d_begin"] or None, b["d_duration"]
            a.d_offsets, a.d_sn_out, a.d_time_out = b["d_offsets"], b["d_sn_out"] or None, b["d_time_out"] or None
            a.hit_capacity, a.d_status = b["capacity"], b["d_status"] or None
        done = ctypes.c_int32(-1)
        code = N.load().hlsmap_get_segment_list_dev_batches(m._h, 9, arr, ctypes.c_void_p(stream.cuda_stream), ctypes.byref(done))
        assert code == N.ERR_INVALID and done.value == 4
        torch.cuda.synchronize()
        for s in sets[:4]:
            assert np.array_equal(s["off"].cpu().numpy().astype(np.int64), s["ref"][0])
        assert int(sets[5]["off"][0].item()) == -1           # nothing behind the broken batch ran
    finally:
        ctx.set_batch_overlap(False)
    m.close()


def test_several_batches_with_one_call(ctx):
    """hlsmap_get_segment_list_dev_batches: the batches of one call behave like so many calls of
    hlsmap_get_segment_list_dev -- each bit-for-bit the oracle's answer, with and without overlap
    between consecutive batches, count-only batches and an empty one among them; a batch that cannot
    be launched stops the call and says how many were issued."""
    import torch
    dev = torch.device("cuda:0")
    t, _ = synth.config2(n_queries=4)
    m = ctx.upload(t)
    stream = torch.cuda.Stream(device=dev)
    sets = []
    for i, nq in enumerate((60_000, 0, 1_000_003, 513, 250_000)):
        _, q = synth.config2(seed=60 + i, n_queries=max(nq, 1))
        q = QueryBatch(q.playlist[:nq], q.level[:nq], q.begin[:nq], q.duration[:nq])
        ref = oracle_list(t, q)
        H = int(ref[0][-1])
        sets.append({"ref": ref, "H": H, "nq": nq,
                     "pl": torch.from_numpy(q.playlist).to(dev), "lv": torch.from_numpy(q.level).to(dev),
                     "bg": torch.from_numpy(q.begin).to(dev), "du": torch.from_numpy(q.duration).to(dev),
                     "off": torch.full((nq + 1,), -1, dtype=torch.int32, device=dev),
                     "sn": torch.empty(max(H, 1), dtype=torch.int32, device=dev),
                     "tm": torch.empty(max(H, 1), dtype=torch.float64, device=dev)})
    batch = lambda s, fill=True: {"nq": s["nq"], "d_playlist": s["pl"].data_ptr(), "d_level": s["lv"].data_ptr(),
                                  "d_begin": s["bg"].data_ptr(), "d_duration": s["du"].data_ptr(),
                                  "d_offsets": s["off"].data_ptr(), "d_sn_out": s["sn"].data_ptr() if fill else 0,
                                  "d_time_out": s["tm"].data_ptr() if fill else 0, "capacity": s["H"] if fill else 0}
    torch.cuda.synchronize()
    for overlap in (False, True):
        ctx.set_batch_overlap(overlap)
        try:
            for s in sets:
                s["off"].fill_(-1)
            torch.cuda.synchronize()
            assert m.get_segment_list_dev_batches([batch(s) for s in sets] * 3, stream.cuda_stream) == 3 * len(sets)
            last = m.last_result(stream.cuda_stream)
            assert last.total_hits == sets[-1]["H"]
            for s in sets:
                assert np.array_equal(s["off"].cpu().numpy().astype(np.int64), s["ref"][0])
                assert np.array_equal(s["sn"][:s["H"]].cpu().numpy(), s["ref"][2])
                assert np.array_equal(s["tm"][:s["H"]].cpu().numpy().view(np.uint64), s["ref"][3].view(np.uint64))
        finally:
            ctx.set_batch_overlap(False)
    # the caller's events, recorded inside the call, bracket the batches on their stream
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    e1.record(stream)
    torch.cuda.synchronize()
    assert m.get_segment_list_dev_batches([batch(s) for s in sets], stream.cuda_stream,
                                          event_begin=e0.cuda_event, event_end=e1.cuda_event) == len(sets)
    torch.cuda.synchronize()
    assert 0.005 < e0.elapsed_time(e1) < 50.0
    # count-only batches size the buffers; a broken batch in the middle stops the call there
    for s in sets:
        s["off"].fill_(-1)
    assert m.get_segment_list_dev_batches([batch(s, fill=False) for s in sets], stream.cuda_stream) == len(sets)
    torch.cuda.synchronize()
    assert [int(s["off"][-1].item()) for s in sets] == [s["H"] for s in sets]
    bad = dict(batch(sets[0]), d_level=0)
    done = ctypes.c_int32(-1)
    arr = (N.Batch * 3)()
    for a, b in zip(arr, (batch(sets[3]), bad, batch(sets[0]))):
        a.nq, a.d_playlist, a.d_level, a.d_begin, a.d_duration = b["nq"], b["d_playlist"], b["d_level"] or None, b["d_begin"], b["d_duration"]
        a.d_offsets, a.d_sn_out, a.d_time_out, a.hit_capacity = b["d_offsets"], b["d_sn_out"], b["d_time_out"], b["capacity"]
    code = N.load().hlsmap_get_segment_list_dev_batches(m._h, 3, arr, ctypes.c_void_p(stream.cuda_stream), ctypes.byref(done))
    assert code == N.ERR_INVALID and done.value == 1
    torch.cuda.synchronize()
    m.close()

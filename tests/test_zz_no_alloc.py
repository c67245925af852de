"""The allocation-free getSegmentList entry point (hlsmap_get_segment_list_noalloc_dev).

It exists for hosts that may not call cudaMalloc while they run (real-time threads, plugins
inside a process that owns all device memory).  Its look-back scratch is a fixed array in the
library image instead of context memory, so this file checks, through the C ABI, that it
gives bit-for-bit what the regular device-pointer entry point gives AND what the CPU oracle
gives (a fault common to both launch paths would otherwise pass) -- on two calls in a row,
because the second call finds the first call's leftovers in that array -- that a batch beyond
the array's capacity is refused with a message instead of overrunning it, and, where the box
has a second GPU, that the array is looked up per device.

Named test_zz_* so pytest collects it last: a failure here cannot hide another test's result
under `-x`.  It makes its own context, so it also passes alone in a fresh process.
"""
import ctypes

import numpy as np
import pytest

import oracle
import hlsjs_p2p_wrapper_b200 as pk
from hlsjs_p2p_wrapper_b200 import synth, _native as N
from hlsjs_p2p_wrapper_b200.tables import FragmentTables, QueryBatch, TRACK_PARSED

pytestmark = pytest.mark.gpu


def _device_batch(torch, dev, q, cap):
    d = {"pl": torch.from_numpy(q.playlist).to(dev), "lv": torch.from_numpy(q.level).to(dev),
         "bg": torch.from_numpy(q.begin).to(dev), "du": torch.from_numpy(q.duration).to(dev)}
    out = lambda: {"off": torch.full((len(q) + 1,), -7, dtype=torch.int32, device=dev),
                   "sn": torch.full((cap,), -7, dtype=torch.int32, device=dev),
                   "tm": torch.full((cap,), -7.0, dtype=torch.float64, device=dev),
                   "st": torch.full((len(q),), 9, dtype=torch.uint8, device=dev)}
    return d, out


def _run(fn, nq, d, o, cap, stream):
    return fn(nq, d["pl"].data_ptr(), d["lv"].data_ptr(), d["bg"].data_ptr(), d["du"].data_ptr(),
              o["off"].data_ptr(), o["sn"].data_ptr(), o["tm"].data_ptr(), cap,
              d_status=o["st"].data_ptr(), stream=stream, sync=True)


def _same(torch, a, b, hits):
    assert torch.equal(a["off"], b["off"]) and torch.equal(a["st"], b["st"])
    assert torch.equal(a["sn"][:hits], b["sn"][:hits])
    assert torch.equal(a["tm"][:hits].view(torch.int64), b["tm"][:hits].view(torch.int64))  # bit patterns


def _same_as_oracle(t, q, o, hits):
    """The entry point's device output against the CPU restatement of media-map.js:27-54."""
    off, status, sn, tm = oracle.get_segment_list_batch(t.playlist_base, t.track_off, t.track_state, t.start,
                                                        t.sn, q.playlist, q.level, q.begin, q.duration)
    assert int(off[-1]) == hits
    assert np.array_equal(o["off"].cpu().numpy().astype(np.int64), off)
    assert np.array_equal(o["st"].cpu().numpy(), status)
    assert np.array_equal(o["sn"][:hits].cpu().numpy(), sn)
    assert np.array_equal(o["tm"][:hits].cpu().numpy().view(np.uint64), tm.view(np.uint64))


@pytest.mark.parametrize("which", ["search_small_map", "search_window_map", "filter"])
def test_noalloc_matches_regular_entry_point_twice_in_a_row(which):
    import torch
    ctx = pk.Context(0)
    dev = torch.device("cuda:0")
    if which == "search_small_map":
        t, q1 = synth.config2(seed=31, n_queries=70_001)
        _, q2 = synth.config2(seed=32, n_queries=33_333)
    elif which == "search_window_map":
        t, q1 = synth.config5(seed=33, n_playlists=800, queries_per_playlist=60)
        q2 = QueryBatch(q1.playlist[:20_000].copy(), q1.level[:20_000].copy(),
                        q1.begin[:20_000].copy(), q1.duration[:20_000].copy())
    else:
        rng = np.random.default_rng(4)
        cols = [rng.integers(0, 400, n) / 4.0 for n in (50, 300, 7)]      # unsorted tables
        off = np.concatenate([[0], np.cumsum([len(c) for c in cols])]).astype(np.int64)
        t = FragmentTables(np.array([0, 3], np.int32), off, np.full(3, TRACK_PARSED, np.uint8),
                           np.full(3, 1, np.int32), np.concatenate(cols),
                           np.arange(int(off[-1]), dtype=np.int32), np.ones(int(off[-1])))
        mk = lambda n: QueryBatch(np.zeros(n, np.int32), rng.integers(-1, 4, n).astype(np.int32),
                                  rng.integers(-20, 420, n) / 4.0, rng.integers(-8, 200, n) / 4.0)
        q1, q2 = mk(9_000), mk(2_500)
    m = ctx.upload(t)
    stream = torch.cuda.Stream(device=dev)
    for q in (q1, q2):                       # two calls in a row through the same fixed scratch
        _, total = m.count_segments(q.playlist, q.level, q.begin, q.duration)
        cap = int(total) + 5
        d, out = _device_batch(torch, dev, q, cap)
        torch.cuda.synchronize()
        ref_o, new_o = out(), out()
        ref = _run(m.get_segment_list_dev, len(q), d, ref_o, cap, stream.cuda_stream)
        new = _run(m.get_segment_list_noalloc_dev, len(q), d, new_o, cap, stream.cuda_stream)
        assert (new.total_hits, new.n_unparsed, new.n_missing, new.path) == \
               (ref.total_hits, ref.n_unparsed, ref.n_missing, ref.path)
        assert new.total_hits == total and new.kernel_launches == 1
        _same(torch, ref_o, new_o, int(total))
        _same_as_oracle(t, q, new_o, int(total))
    m.close()
    ctx.close()


def test_noalloc_on_a_second_gpu_uses_that_gpus_array():
    """The fixed scratch array exists once per GPU and its address is looked up per device: a call
    on cuda:1 after a call on cuda:0 (and back) must use its own copy.  Needs two visible GPUs."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU visible: the per-device look-up has nothing to tell apart")
    t, q = synth.config2(seed=36, n_queries=40_003)
    ctxs = [pk.Context(0), pk.Context(1)]
    maps = [c.upload(t) for c in ctxs]
    _, total = maps[0].count_segments(q.playlist, q.level, q.begin, q.duration)
    cap = int(total) + 3
    for which in (0, 1, 0, 1):
        dev = torch.device("cuda", which)
        with torch.cuda.device(dev):
            d, out = _device_batch(torch, dev, q, cap)
            o = out()
            stream = torch.cuda.Stream(device=dev)
            torch.cuda.synchronize(dev)
            res = _run(maps[which].get_segment_list_noalloc_dev, len(q), d, o, cap, stream.cuda_stream)
            assert res.total_hits == total
            _same_as_oracle(t, q, o, int(total))
    for m in maps:
        m.close()
    for c in ctxs:
        c.close()


def test_noalloc_refuses_batches_beyond_its_fixed_scratch():
    import torch
    ctx = pk.Context(0)
    t, q = synth.config2(seed=35, n_queries=64)
    m = ctx.upload(t)
    dev = torch.device("cuda:0")
    d, out = _device_batch(torch, dev, q, 16)
    o = out()
    too_many = 65536 * 4096 + 1             # more than 65536 tiles whatever the tile size (<= 4096)
    with pytest.raises(N.HlsMapError) as e:
        # the size check comes before any pointer is touched, so small buffers are fine here
        m.get_segment_list_noalloc_dev(too_many, d["pl"].data_ptr(), d["lv"].data_ptr(),
                                       d["bg"].data_ptr(), d["du"].data_ptr(), o["off"].data_ptr(),
                                       o["sn"].data_ptr(), o["tm"].data_ptr(), 16)
    assert e.value.code == N.ERR_INVALID
    assert "65536" in e.value.message and "split the batch" in e.value.message
    # an empty batch and a capacity shortfall behave like the regular entry point
    res = m.get_segment_list_noalloc_dev(0, 0, d["lv"].data_ptr(), d["bg"].data_ptr(), d["du"].data_ptr(),
                                         o["off"].data_ptr(), o["sn"].data_ptr(), o["tm"].data_ptr(), 16)
    assert res.total_hits == 0 and int(o["off"][0].item()) == 0
    with pytest.raises(N.HlsMapError) as e:
        m.get_segment_list_noalloc_dev(len(q), d["pl"].data_ptr(), d["lv"].data_ptr(), d["bg"].data_ptr(),
                                       d["du"].data_ptr(), o["off"].data_ptr(), o["sn"].data_ptr(),
                                       o["tm"].data_ptr(), 1)
    assert e.value.code == N.ERR_CAPACITY
    m.close()
    ctx.close()

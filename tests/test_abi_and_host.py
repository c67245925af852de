"""CPU-only checks: the C-ABI library loads and exports every symbol include/hlsmap.h
declares (no compute without a GPU), host-side table conversion, generators, sharding."""
import os
import re

import numpy as np
import pytest

import oracle
from hlsjs_p2p_wrapper_b200 import _native as N, FragmentTables, QueryBatch, synth, shard
from hlsjs_p2p_wrapper_b200 import build as B

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_builds_loads_and_exports_every_declared_symbol():
    B.build()
    lib = N.load()
    header = open(os.path.join(ROOT, "include", "hlsmap.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(hlsmap_[a-z0-9_]+)\s*\(", header))
    assert declared == set(N.SYMBOLS), declared ^ set(N.SYMBOLS)
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.hlsmap_version() == 200
    assert isinstance(lib.hlsmap_last_error(), bytes)


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import hlsjs_p2p_wrapper_b200 as pk
    with pytest.raises(N.HlsMapError) as e:
        pk.Context(0)
    assert e.value.code == N.ERR_CUDA


def test_sass_has_bulk_async_copies():
    """The staging path is TMA-class bulk copy (UBLKCP) + mbarrier, compiled for sm_100a."""
    import shutil
    import subprocess
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    B.build()
    out = subprocess.check_output(["cuobjdump", "-sass", B.LIB]).decode()
    assert "sm_100a" in out and "UBLKCP" in out and "SYNCS" in out


def _mulberry32_scalar(seed, n):
    a, out = seed & 0xFFFFFFFF, []
    for _ in range(n):
        a = (a + 0x6D2B79F5) & 0xFFFFFFFF
        t = ((a ^ (a >> 15)) * (a | 1)) & 0xFFFFFFFF
        t = (((t + (((t ^ (t >> 7)) * (t | 61)) & 0xFFFFFFFF)) & 0xFFFFFFFF) ^ t)
        out.append((t ^ (t >> 14)) & 0xFFFFFFFF)
    return out


def test_mulberry32_vectorised_equals_scalar_and_streams_continue():
    for seed in (0, 1, 0xDEADBEEF, 2**32 - 1):
        r = synth.Mulberry32(seed)
        got = list(map(int, r.u32(50))) + list(map(int, r.u32(77)))
        assert got == _mulberry32_scalar(seed, 127)


def test_config_generators_are_deterministic_and_well_formed():
    t1, q1 = synth.config2(n_queries=2000)
    t2, q2 = synth.config2(n_queries=2000)
    assert np.array_equal(t1.start, t2.start) and np.array_equal(q1.begin.view(np.uint64), q2.begin.view(np.uint64))
    assert t1.n_tracks == 3 and t1.n_fragments == 3000
    assert all(oracle.is_monotone(t1.start[t1.track_off[i]:t1.track_off[i + 1]]) for i in range(3))
    d = np.diff(t1.start[:1000])
    assert d.min() >= 2.0 and d.max() <= 10.0 and np.array_equal(d * 1024, np.round(d * 1024))
    assert np.isnan(q1.begin).any() or np.isnan(q1.duration).any()       # adversarial mix present
    t3, q3 = synth.config3(n_players=50)
    assert t3.n_tracks == 150 and (np.diff(t3.track_off) == 6).all()
    assert all(oracle.is_monotone(t3.start[o:o + 6]) for o in t3.track_off[:-1])
    assert t3.sn.min() >= 0
    t5, q5 = synth.config5(n_playlists=40, queries_per_playlist=10)
    lens = np.diff(t5.track_off)
    assert lens.min() >= 6 and lens.max() <= 1024 and len(q5) == 400
    assert (np.diff(q5.playlist) >= 0).all()
    t4, v, o = synth.config4(n_views=5000)
    assert t4.n_tracks == 8 and t4.n_fragments == 80000
    assert (v["sn"] != o["sn"]).any() or (v["level"] != o["level"]).any()


def test_tables_from_hls_levels_and_slicing():
    lv = lambda fr: {"url": ["a", "b"], "details": {"fragments": fr}}
    f = lambda n: [{"sn": 5 + i, "start": 2.0 * i, "duration": 2.0} for i in range(n)]
    t = FragmentTables.from_hls_levels([[lv(f(3)), {"url": ["a"]}], None, [None, lv(f(2))]])
    assert list(t.playlist_base) == [0, 2, 2, 4]
    assert list(t.track_state) == [2, 1, 0, 2] and list(t.track_off) == [0, 3, 3, 3, 5]
    assert list(t.url_count) == [2, 1, 0, 2]
    s = t.slice_playlists(2, 3)
    assert list(s.playlist_base) == [0, 2] and list(s.track_off) == [0, 0, 2] and list(s.sn) == [5, 6]
    assert np.array_equal(t.dur0(), [2.0, np.nan, np.nan, 2.0], equal_nan=True)


def _oracle_list(t, q):
    return oracle.get_segment_list_batch(t.playlist_base, t.track_off, t.track_state, t.start, t.sn,
                                         q.playlist, q.level, q.begin, q.duration)


@pytest.mark.parametrize("n_shards", [1, 2, 3, 8])
def test_shard_split_merge_roundtrip(n_shards):
    t, q = synth.config5(seed=9, n_playlists=64, queries_per_playlist=20)
    perm = np.random.default_rng(1).permutation(len(q))
    q = QueryBatch(q.playlist[perm], q.level[perm], q.begin[perm], q.duration[perm])
    q.playlist[:5] = [-1, 64, 1000, 0, 63]                  # out-of-range playlists too
    bounds = shard.plan(t, q, n_shards)
    assert bounds[0] == 0 and bounds[-1] == 64 and (np.diff(bounds) >= 0).all()
    shards, order = shard.split(t, q, bounds)
    res = [_oracle_list(ts, qs) for ts, qs in shards]
    off, status, sn, tm = shard.merge(res, order)
    eo, es, esn, etm = _oracle_list(t, q)
    assert np.array_equal(off, eo) and np.array_equal(status, es)
    assert np.array_equal(sn, esn) and np.array_equal(tm.view(np.uint64), etm.view(np.uint64))

"""world_size-2 `gloo` test of the N > 1 path on CPU: each rank resolves its playlist shard
(the oracle stands in for the GPU here), a host-side gather rebuilds the global answer, and
the timing reduction bench.py uses (max over ranks) works.  No data-path collective."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from hlsjs_p2p_wrapper_b200 import QueryBatch, synth, shard


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    t, q = synth.config5(seed=12, n_playlists=48, queries_per_playlist=25)
    bounds = shard.plan(t, q, world)
    shards, order = shard.split(t, q, bounds)
    ts, qs = shards[rank]
    mine = oracle.get_segment_list_batch(ts.playlist_base, ts.track_off, ts.track_state, ts.start,
                                         ts.sn, qs.playlist, qs.level, qs.begin, qs.duration)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)                  # host-side gather only
    elapsed = torch.tensor([1.0 + rank], dtype=torch.float64)
    dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)           # bench.py: max over ranks
    dist.barrier()
    if rank == 0:
        off, status, sn, tm = shard.merge(gathered, order)
        eo, es, esn, etm = oracle.get_segment_list_batch(
            t.playlist_base, t.track_off, t.track_state, t.start, t.sn,
            q.playlist, q.level, q.begin, q.duration)
        ok = (np.array_equal(off, eo) and np.array_equal(status, es) and np.array_equal(sn, esn)
              and np.array_equal(tm.view(np.uint64), etm.view(np.uint64))
              and float(elapsed.item()) == float(world))
        with open(out_path, "w") as f:
            f.write("ok" if ok else "mismatch")
    dist.destroy_process_group()


def test_two_rank_shard_gather(tmp_path):
    out = str(tmp_path / "result.txt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert open(out).read() == "ok"


def _text_worker(rank, world, port, out_path):
    """Playlists still in text form: every rank parses its own shard of the text (oracle in place of the
    GPU ingest) and answers the queries routed to it; rank 0 gathers and compares with the unsharded run."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    base, state, texts = synth.m3u8_corpus(seed=21, n_playlists=60, unparsed_every=11)
    off = np.concatenate([[0], np.cumsum([len(t) for t in texts])]).astype(np.int64)
    text = b"".join(texts)
    bounds = np.asarray([0, 27, 60])
    b, st, tx, to = shard.split_texts(base, state, text, off, bounds)[rank]
    mine_tab = oracle.parse_m3u8(tx, to, st)
    rng = np.random.default_rng(5)
    n = 4000
    q = QueryBatch(rng.integers(0, 60, n).astype(np.int32), rng.integers(0, 3, n).astype(np.int32),
                   rng.uniform(-5, 80, n), rng.uniform(0, 40, n))
    whole = oracle.parse_m3u8(text, off, state)
    from hlsjs_p2p_wrapper_b200.tables import FragmentTables
    routing = FragmentTables(base, whole["track_off"], state, np.ones(len(state), np.int32), whole["start"],
                             whole["sn"], whole["duration"])
    shards, order = shard.split(routing, q, bounds)
    qs = shards[rank][1]
    mine = oracle.get_segment_list_batch(b, mine_tab["track_off"], st, mine_tab["start"], mine_tab["sn"],
                                         qs.playlist, qs.level, qs.begin, qs.duration)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    if rank == 0:
        o, s, sn, tm = shard.merge(gathered, order)
        eo, es, esn, etm = oracle.get_segment_list_batch(base, whole["track_off"], state, whole["start"], whole["sn"],
                                                         q.playlist, q.level, q.begin, q.duration)
        ok = (np.array_equal(o, eo) and np.array_equal(s, es) and np.array_equal(sn, esn)
              and np.array_equal(tm.view(np.uint64), etm.view(np.uint64)) and len(sn) > 500)
        with open(out_path, "w") as f:
            f.write("ok" if ok else "mismatch")
    dist.destroy_process_group()


def test_two_rank_text_shards(tmp_path):
    out = str(tmp_path / "result.txt")
    mp.spawn(_text_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert open(out).read() == "ok"

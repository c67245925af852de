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

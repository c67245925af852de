#!/usr/bin/env python
"""BASELINE.json configs[4]: getSegmentList over many synthetic playlists, queries grouped by
playlist, checked bit-for-bit against the CPU oracle chunk by chunk (hard
part 9: compare chunk-wise so a mismatch can be localised).

    python scripts/run_config5.py --playlists 1000000 --chunk 50000 --out gpurun_out/config5.json

Each chunk of playlists is its own uploaded map (playlists are independent, so this is the same
sharding-by-playlist the multi-GPU path uses); per chunk: generate, upload, resolve on the GPU
through the C ABI, resolve on the CPU oracle, compare counts, sn and fp64 bit patterns of time.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import oracle
import hlsjs_p2p_wrapper_b200 as pk
from hlsjs_p2p_wrapper_b200 import synth


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--playlists", type=int, default=1_000_000)
    ap.add_argument("--chunk", type=int, default=50_000)
    ap.add_argument("--queries-per-playlist", type=int, default=100)
    ap.add_argument("--threads", type=int, default=len(os.sched_getaffinity(0)))
    ap.add_argument("--out", default="gpurun_out/config5.json")
    args = ap.parse_args()

    ctx = pk.Context(0)
    dev = torch.device("cuda:0")
    stream = torch.cuda.Stream(device=dev)
    tot = {"queries": 0, "hits": 0, "fragments": 0, "mismatched_queries": 0, "mismatched_hits": 0,
           "kernel_ms": 0.0, "alg_bytes": 0.0, "oracle_s": 0.0, "chunks": 0}
    t_all = time.time()
    for c0 in range(0, args.playlists, args.chunk):
        n = min(args.chunk, args.playlists - c0)
        t, q = synth.config5(seed=1000 + c0 // args.chunk, n_playlists=n,
                             queries_per_playlist=args.queries_per_playlist)
        m = ctx.upload(t)
        nq = len(q)
        # device-resident timing of the kernel for this chunk
        d = {k: torch.from_numpy(getattr(q, k)).to(dev) for k in ("playlist", "level", "begin", "duration")}
        _, total = m.count_segments(q.playlist, q.level, q.begin, q.duration)
        off = torch.empty(nq + 1, dtype=torch.int32, device=dev)
        sn = torch.empty(total, dtype=torch.int32, device=dev)
        tm = torch.empty(total, dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        with torch.cuda.stream(stream):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for i in range(3):
                if i == 2:
                    e0.record()
                m.get_segment_list_dev(nq, d["playlist"].data_ptr(), d["level"].data_ptr(),
                                       d["begin"].data_ptr(), d["duration"].data_ptr(), off.data_ptr(),
                                       sn.data_ptr(), tm.data_ptr(), total, stream=stream.cuda_stream)
            e1.record()
        torch.cuda.synchronize()
        kms = e0.elapsed_time(e1)
        g_off, g_sn, g_tm = off.cpu().numpy().astype(np.int64), sn.cpu().numpy(), tm.cpu().numpy()
        t0 = time.time()
        o_off, o_st, o_sn, o_tm = oracle.get_segment_list_batch(
            t.playlist_base, t.track_off, t.track_state, t.start, t.sn,
            q.playlist, q.level, q.begin, q.duration, threads=args.threads)
        tot["oracle_s"] += time.time() - t0
        bad_q = int((np.diff(g_off) != np.diff(o_off)).sum())
        if bad_q == 0:
            bad_h = int((g_sn != o_sn).sum() + (g_tm.view(np.uint64) != o_tm.view(np.uint64)).sum())
        else:
            bad_h = -1
        tot["queries"] += nq
        tot["hits"] += int(total)
        tot["fragments"] += t.n_fragments
        tot["mismatched_queries"] += bad_q
        tot["mismatched_hits"] += max(bad_h, 0) if bad_q == 0 else int(total)
        tot["kernel_ms"] += kms
        tot["alg_bytes"] += 28.0 * nq + 12.0 * total + 12.0 * t.n_fragments
        tot["chunks"] += 1
        info = m.info()
        m.close()
        del d, off, sn, tm
        print("chunk %3d: %d playlists %d frags %d queries %d hits  kernel %.3f ms  mismatches q=%d h=%d  stage_all=%d"
              % (tot["chunks"], n, t.n_fragments, nq, total, kms, bad_q, bad_h, info["stage_all"]), flush=True)
    tot["wall_s"] = time.time() - t_all
    tot["queries_per_s_kernel"] = tot["queries"] / (tot["kernel_ms"] * 1e-3)
    tot["achieved_GBs"] = tot["alg_bytes"] / (tot["kernel_ms"] * 1e-3) / 1e9
    peaks = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    peak = json.load(open(peaks))["hbm_gbs"] if os.path.exists(peaks) else 6650.0
    tot["roofline_frac"] = tot["achieved_GBs"] / peak
    tot["hits_per_query"] = tot["hits"] / max(tot["queries"], 1)
    tot["config"] = vars(args)
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    json.dump(tot, open(args.out, "w"), indent=1)
    print(json.dumps(tot))
    return 0 if tot["mismatched_queries"] == 0 and tot["mismatched_hits"] == 0 else 1


if __name__ == "__main__":
    sys.exit(main())

"""Times playlist ingest (hlsmap_parse_m3u8_dev) on text already resident in HBM and checks a sample
against the oracle.  Usage: python scripts/run_ingest.py [out.json]"""
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle                                                    # noqa: E402  (checker + CPU baseline)
from hlsjs_p2p_wrapper_b200 import synth, _native as N           # noqa: E402
from hlsjs_p2p_wrapper_b200.engine import Context                # noqa: E402


def run(ctx, name, text, off, reps=5):
    lib = N.load()
    dev = torch.device("cuda", 0)
    d_text, d_off = torch.from_numpy(text).to(dev), torch.from_numpy(off).to(dev)
    t = len(off) - 1
    exp = oracle.parse_m3u8(text, off, threads=os.cpu_count() or 1)
    cap = int(exp["track_off"][-1])
    d_toff = torch.empty(t + 1, dtype=torch.int64, device=dev)
    d_live = torch.empty(t, dtype=torch.uint8, device=dev)
    d_start = torch.empty(cap, dtype=torch.float64, device=dev)
    d_sn = torch.empty(cap, dtype=torch.int32, device=dev)
    d_dur = torch.empty(cap, dtype=torch.float64, device=dev)
    res = N.IngestResult()
    stream = torch.cuda.Stream()
    v = lambda x: ctypes.c_void_p(x.data_ptr())
    times = []
    for i in range(reps + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(stream)
        N.check(lib.hlsmap_parse_m3u8_dev(ctx._h, t, None, v(d_text), v(d_off), len(text), v(d_toff), v(d_live),
                                          v(d_start), v(d_sn), v(d_dur), cap, ctypes.byref(res),
                                          ctypes.c_void_p(stream.cuda_stream)))
        e1.record(stream)
        torch.cuda.synchronize()
        if i >= 2:
            times.append(e0.elapsed_time(e1))
    ok = (np.array_equal(d_toff.cpu().numpy(), exp["track_off"]) and np.array_equal(d_sn.cpu().numpy(), exp["sn"])
          and np.array_equal(d_start.cpu().numpy().view(np.int64), exp["start"].view(np.int64))
          and np.array_equal(d_dur.cpu().numpy().view(np.int64), exp["duration"].view(np.int64))
          and np.array_equal(d_live.cpu().numpy(), exp["live"]))
    t0 = time.perf_counter()
    oracle.parse_m3u8(text, off, threads=os.cpu_count() or 1)
    cpu = time.perf_counter() - t0
    ms = float(np.median(times))
    # algorithmic bytes: the text once, 8 B duration written and read, 8 B start + 4 B sn written
    alg = len(text) + cap * (8 + 8 + 8 + 4)
    return dict(case=name, text_MB=len(text) / 1e6, tracks=t, fragments=cap, parity=bool(ok), ms_whole_call=ms,
                text_GBps=len(text) / ms / 1e6, algorithmic_GBps=alg / ms / 1e6, fragments_per_s=cap / ms * 1e3,
                cpu_oracle_s=cpu, cpu_threads=os.cpu_count(), cpu_fragments_per_s=cap / cpu,
                kernel_launches=res.kernel_launches)


def main():
    ctx = Context(0)
    out = []
    text, off = synth.m3u8_block(n_tracks=8000, n_frag=800)              # VOD: 8000 x 800 fragments
    out.append(run(ctx, "vod 8000 playlists x 800 fragments", text, off))
    text, off = synth.m3u8_block(n_tracks=1_000_000, n_frag=6)           # live: 1e6 windows of 6
    out.append(run(ctx, "live 1e6 playlists x 6 fragments", text, off))
    for o in out:
        print(json.dumps(o))
    if len(sys.argv) > 1:
        json.dump(out, open(sys.argv[1], "w"), indent=1)


if __name__ == "__main__":
    main()

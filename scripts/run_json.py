"""Times the JSON form of views on the GPU (hlsmap_view_to_json_dev / hlsmap_view_from_json_dev, columns
resident in HBM) and checks a sample against the oracle.  Usage: python scripts/run_json.py [out.json]"""
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle                                                    # noqa: E402  (checker + CPU baseline)
from hlsjs_p2p_wrapper_b200 import _native as N                  # noqa: E402
from hlsjs_p2p_wrapper_b200.engine import Context                # noqa: E402


def main():
    ctx, lib, dev = Context(0), N.load(), torch.device("cuda", 0)
    n = 8_000_000
    rng = np.random.default_rng(1)
    lv, ui = rng.integers(0, 8, n).astype(np.int32), rng.integers(0, 2, n).astype(np.int32)
    sn = rng.integers(0, 2 ** 31 - 1, n).astype(np.int32)
    tm = np.cumsum(rng.integers(2000, 10001, n) / 1000.0) % 86400.0      # playlist times, ms steps, fp64 sums
    d = [torch.from_numpy(a).to(dev) for a in (lv, ui, sn, tm)]
    d_off = torch.empty(n + 1, dtype=torch.int32, device=dev)
    d_chars = torch.empty(104 * n, dtype=torch.uint8, device=dev)
    stream = torch.cuda.Stream()
    v = lambda x: ctypes.c_void_p(x.data_ptr())
    s = ctypes.c_void_p(stream.cuda_stream)
    total = ctypes.c_int64(0)

    def timed(fn, reps=5):
        ts = []
        for i in range(reps + 2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record(stream)
            fn()
            e1.record(stream)
            torch.cuda.synchronize()
            if i >= 2:
                ts.append(e0.elapsed_time(e1))
        return float(np.median(ts))

    ms_w = timed(lambda: N.check(lib.hlsmap_view_to_json_dev(ctx._h, n, v(d[0]), v(d[1]), v(d[2]), v(d[3]), v(d_off),
                                                             v(d_chars), 104 * n, s, None)))
    N.check(lib.hlsmap_view_to_json_dev(ctx._h, n, v(d[0]), v(d[1]), v(d[2]), v(d[3]), v(d_off), v(d_chars),
                                        104 * n, s, ctypes.byref(total)))
    nchar = total.value
    o = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(3)]
    o_t = torch.empty(n, dtype=torch.float64, device=dev)
    o_f, o_s = torch.empty(n, dtype=torch.uint8, device=dev), torch.empty(n, dtype=torch.uint8, device=dev)
    ms_r = timed(lambda: N.check(lib.hlsmap_view_from_json_dev(ctx._h, n, v(d_chars), v(d_off), v(o[0]), v(o[1]), v(o[2]),
                                                               v(o_t), v(o_f), v(o_s), s)))
    ok_back = (torch.equal(o[0], d[0]) and torch.equal(o[1], d[1]) and torch.equal(o[2], d[2])
               and torch.equal(o_t, d[3]) and int(o_s.sum()) == 0)
    m = 400_000                                                          # oracle on a sample
    t0 = time.perf_counter()
    eoff, echars = oracle.view_to_json(lv[:m], ui[:m], sn[:m], tm[:m])
    cpu_w = time.perf_counter() - t0
    off = d_off[:m + 1].cpu().numpy()
    ok_text = np.array_equal(off, eoff) and np.array_equal(d_chars[:off[m]].cpu().numpy(), echars)
    t0 = time.perf_counter()
    oracle.view_from_json(echars, eoff)
    cpu_r = time.perf_counter() - t0
    # algorithmic bytes: 20 B of columns + 4 B offset + the characters, either way
    alg = n * 24 + nchar
    out = dict(views=n, chars=nchar, chars_per_view=nchar / n, parity_text=bool(ok_text), roundtrip_exact=bool(ok_back),
               stringify_ms=ms_w, stringify_views_per_s=n / ms_w * 1e3, stringify_GBps=alg / ms_w / 1e6,
               parse_ms=ms_r, parse_views_per_s=n / ms_r * 1e3, parse_GBps=(alg + 2 * n) / ms_r / 1e6,
               cpu_oracle_stringify_views_per_s=m / cpu_w, cpu_oracle_parse_views_per_s=m / cpu_r, cpu_threads=1)
    print(json.dumps(out))
    if len(sys.argv) > 1:
        json.dump(out, open(sys.argv[1], "w"), indent=1)


if __name__ == "__main__":
    main()

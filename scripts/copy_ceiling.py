#!/usr/bin/env python
"""What the host <-> GPU copy path of this box sustains with 1, 2, 4, 8 GPUs copying at once: the
ceiling bench.py's `e2e` (host pointers in, host pointers out) can reach.

Per GPU and step, the traffic of one configs[1] batch in list form: 24 MB of queries up
(4 pinned arrays) and 157 MB of offsets + hits down (3 pinned arrays), the two directions on
their own streams so that they overlap (PCIe is full duplex), as the library's chunked pipeline
does.  One host thread per GPU; plain cudaMemcpyAsync through torch, nothing of libhlsmap.

    python scripts/copy_ceiling.py [--steps 12] [--out profiles/copy_ceiling_r2.json]
"""
import argparse
import json
import os
import threading
import time

import torch

UP = [4_000_000, 4_000_000, 8_000_000, 8_000_000]        # playlist, level, begin, duration (bytes)
DOWN = [4_000_004, 51_070_332, 102_140_664]              # offsets, sn, time of 12.77e6 hits


def worker(dev, steps, barrier, out):
    torch.cuda.set_device(dev)
    s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    h_up = [torch.empty(n, dtype=torch.uint8).pin_memory() for n in UP]
    h_dn = [torch.empty(n, dtype=torch.uint8).pin_memory() for n in DOWN]
    d_up = [torch.empty(n, dtype=torch.uint8, device=dev) for n in UP]
    d_dn = [torch.empty(n, dtype=torch.uint8, device=dev) for n in DOWN]

    def step():
        with torch.cuda.stream(s_up):
            for h, d in zip(h_up, d_up):
                d.copy_(h, non_blocking=True)
        with torch.cuda.stream(s_dn):
            for h, d in zip(h_dn, d_dn):
                h.copy_(d, non_blocking=True)
    for _ in range(3):
        step()
    torch.cuda.synchronize(dev)
    barrier.wait()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize(dev)
    out[dev] = time.perf_counter() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    n_dev = torch.cuda.device_count()
    rows = []
    for n in (1, 2, 4, 8):
        if n > n_dev:
            break
        out = {}
        barrier = threading.Barrier(n)
        ths = [threading.Thread(target=worker, args=(d, args.steps, barrier, out)) for d in range(n)]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        worst = max(out.values())
        per_step = sum(UP) + sum(DOWN)
        rows.append({"gpus": n, "s_per_step_max_over_gpus": worst / args.steps,
                     "GBps_per_gpu": per_step / (worst / args.steps) / 1e9,
                     "GBps_aggregate": n * per_step / (worst / args.steps) / 1e9,
                     "steps_per_s_aggregate": n * args.steps / worst,
                     "equivalent_e2e_queries_per_s": n * args.steps * 1_000_000 / worst})
        print(json.dumps(rows[-1]), flush=True)
    rec = {"what": "concurrent pinned cudaMemcpyAsync, 24 MB up + 157 MB down per GPU and step on two streams",
           "host_threads": len(os.sched_getaffinity(0)), "rows": rows}
    if args.out:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        json.dump(rec, open(args.out, "w"), indent=1)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())

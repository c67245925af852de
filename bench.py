#!/usr/bin/env python
"""bench.py -- getSegmentList queries/s (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one batched getSegmentList pass (list form) over BASELINE.json configs[1]: one VOD
playlist of 3 levels x 1000 fragments and 10^6 random (playlist, level, beginTime, duration)
queries.  With N > 1 every rank owns its own playlist shard and query batch (weak scaling, no
collective on the data path); `value` is total queries / max-over-ranks device time.

  value     inputs resident in HBM, K launches timed with CUDA events on the launch stream
  e2e       the same step through the C ABI's HOST-pointer entry point on pinned host
            buffers: H2D of the queries and D2H of offsets + hits are inside the timed region
  roofline  algorithmic bytes (28 B/query + 12 B/hit + the tables once) / mean launch time,
            against the measured HBM copy peak of MEASURED_PEAKS.json
  cpu_baseline  the C restatement of media-map.js:42-51 (oracle/) on the host cores, timed on
            rank 0 -- "port": no JavaScript engine exists on these boxes

--impl reference times that CPU restatement alone with every host thread (the reference is
JavaScript) and prints the same line with "impl": "reference".
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "getSegmentList queries/s"
UNIT = "queries/s"
L2_BYTES = 126 * 1024 * 1024


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--queries", type=int, default=1_000_000)
    ap.add_argument("--e2e-steps", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-rows", action="store_true",
                    help="skip the short timings of the rows either side of the path (ingest, JSON)")
    ap.add_argument("--no-sharded", action="store_true",
                    help="skip the BASELINE.json configs[4] leg (10^8 queries sharded by playlist, ~2 min at N=1)")
    ap.add_argument("--config5-chunks", type=int, default=20,
                    help="corpus chunks of 50 000 playlists for the sharded leg (20 = the full 10^6 playlists)")
    return ap.parse_args()


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def node_version():
    exe = shutil.which("node") or shutil.which("nodejs")
    if not exe:
        return None
    try:
        return subprocess.check_output([exe, "--version"], timeout=10).decode().strip()
    except Exception:
        return None


def workload_config(args, world):
    return {"workload": "BASELINE.json configs[1]: VOD playlist, 3 levels x 1000 fragments, "
                        "%d random (playlist, level, beginTime, duration) queries per GPU, list form"
                        % args.queries,
            "queries_per_step_per_gpu": args.queries, "levels": 3, "fragments_per_level": 1000,
            "sharding": "by playlist, one shard per GPU, no collective" if world > 1 else "single GPU"}


# ---------------------------------------------------------------------------------------------
#  CPU arm: the restated reference filter on the host cores
# ---------------------------------------------------------------------------------------------
def cpu_filter_rate(tables, q, threads, min_seconds, max_reps=50):
    import oracle
    oracle.build()
    reps, t0 = 0, time.perf_counter()
    hits = 0
    while True:
        hits, _ = oracle.bench_get_segment_list(tables.playlist_base, tables.track_off,
                                                tables.track_state, tables.start, tables.sn,
                                                q.playlist, q.level, q.begin, q.duration, threads)
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= min_seconds or reps >= max_reps:
            break
    return reps * len(q) / dt, reps, dt, hits


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return 0
    from hlsjs_p2p_wrapper_b200 import synth
    import oracle
    oracle.build()
    tables, q = synth.config2(seed=2, n_queries=args.queries)
    threads = host_threads()

    def step():
        return oracle.bench_get_segment_list(tables.playlist_base, tables.track_off, tables.track_state,
                                             tables.start, tables.sn, q.playlist, q.level, q.begin,
                                             q.duration, threads)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        hits, _ = step()
    dt = time.perf_counter() - t0
    value = args.steps * len(q) / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(args, world),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "every step = the full %d-query batch through the C restatement of "
                                   "media-map.js:42-51 (linear filter + list push), %d pthreads; the "
                                   "reference JS itself cannot run: node is %s"
                                   % (len(q), threads, node_version() or "absent")},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "hits_per_query": hits / len(q), "node": node_version(),
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------
#  clocks sampling during the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock and throttle reasons through NVML every ~0.5 ms while the timed region
    runs (nvidia-smi is far too slow to land a sample inside a region of a millisecond; the
    driver's default 20-step run times 0.75 ms)."""

    def __init__(self, uuid=None, index=0):
        self.uuid, self.index = uuid, index
        self.samples = []          # (t, sm_mhz, reasons_bitmask)
        self.window = None         # (t0, t1) of the timed region
        self._stop = threading.Event()
        self._t = None
        self.max_mhz = None
        self.err = None

    def _run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = None
            if self.uuid:
                for u in (self.uuid, "GPU-" + self.uuid):
                    try:
                        h = nv.nvmlDeviceGetHandleByUUID(u)
                        break
                    except Exception:
                        h = None
            if h is None:
                h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                nv.nvmlDeviceGetCurrentClocksThrottleReasons
            while not self._stop.is_set():
                self.samples.append((time.perf_counter(),
                                     float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)),
                                     int(get_reasons(h))))
                self._stop.wait(0.0005)
        except Exception as e:          # no NVML: report that instead of inventing clocks
            self.err = repr(e)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def wait_first(self, timeout):
        """Block until NVML has delivered a first sample (or failed): a timed region of a
        millisecond then still has the clocks read immediately before and after it."""
        t0 = time.perf_counter()
        while not self.samples and self.err is None and time.perf_counter() - t0 < timeout:
            time.sleep(0.01)

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=10)
        bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20,
                "hw_thermal_slowdown": 0x40, "hw_power_brake_slowdown": 0x80}
        t0, t1 = self.window or (0.0, float("inf"))
        inside = [s for s in self.samples if t0 <= s[0] <= t1]
        if not inside:       # none landed inside: the last sample before and the first after the region
            before = [s for s in self.samples if s[0] < t0][-1:]
            after = [s for s in self.samples if s[0] > t1][:1]
            inside = before + after or self.samples
        sm = [s[1] for s in inside]
        mask = 0
        for s in inside:
            mask |= s[2]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(k for k, b in bits.items() if mask & b),
                "samples_in_timed_region": len([s for s in self.samples if t0 <= s[0] <= t1]),
                "samples": len(self.samples), "error": self.err}


# ---------------------------------------------------------------------------------------------
#  GPU arm
# ---------------------------------------------------------------------------------------------
def other_rows(ctx, torch, peak):
    """Short device-timed runs of the SURVEY 8 'next' rows that sit either side of the path: playlist
    text -> tables (f2) and views -> JSON -> views (f3).  Inputs resident in HBM; parity of both is the
    business of tests/test_gpu_ingest.py and tests/test_gpu_json.py, not of this function."""
    import ctypes
    from hlsjs_p2p_wrapper_b200 import _native as N, synth
    lib, dev = N.load(), torch.device("cuda", torch.cuda.current_device())
    stream = torch.cuda.Stream()
    sp = ctypes.c_void_p(stream.cuda_stream)
    v = lambda x: ctypes.c_void_p(x.data_ptr())

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

    out = {}
    text, off = synth.m3u8_block(n_tracks=6000, n_frag=800)                  # 149 MB of M3U8: > L2
    t, cap = len(off) - 1, 6000 * 800
    d_text, d_off = torch.from_numpy(text).to(dev), torch.from_numpy(off).to(dev)
    d_toff = torch.empty(t + 1, dtype=torch.int64, device=dev)
    d_live = torch.empty(t, dtype=torch.uint8, device=dev)
    d_start, d_dur = (torch.empty(cap, dtype=torch.float64, device=dev) for _ in range(2))
    d_sn = torch.empty(cap, dtype=torch.int32, device=dev)
    res = N.IngestResult()
    ms = timed(lambda: N.check(lib.hlsmap_parse_m3u8_dev(ctx._h, t, None, v(d_text), v(d_off), len(text), v(d_toff),
                                                         v(d_live), v(d_start), v(d_sn), v(d_dur), cap,
                                                         ctypes.byref(res), sp)))
    alg = len(text) + 28 * int(res.n_fragments)
    out["ingest_m3u8"] = {"text_MB": len(text) / 1e6, "fragments": int(res.n_fragments), "ms_per_call": ms,
                          "fragments_per_s": res.n_fragments / ms * 1e3, "text_GBps": len(text) / ms / 1e6,
                          "roofline_frac": alg / ms / 1e6 / peak, "kernel_launches": int(res.kernel_launches)}
    del d_text, d_start, d_dur, d_sn
    n = 4_000_000
    rng = np.random.default_rng(1)
    cols = [torch.from_numpy(a).to(dev) for a in (rng.integers(0, 8, n).astype(np.int32),
                                                  rng.integers(0, 2, n).astype(np.int32),
                                                  rng.integers(0, 2 ** 31 - 1, n).astype(np.int32),
                                                  np.cumsum(rng.integers(2000, 10001, n) / 1000.0) % 86400.0)]
    d_o = torch.empty(n + 1, dtype=torch.int32, device=dev)
    d_c = torch.empty(104 * n, dtype=torch.uint8, device=dev)
    total = ctypes.c_int64(0)
    ms_w = timed(lambda: N.check(lib.hlsmap_view_to_json_dev(ctx._h, n, v(cols[0]), v(cols[1]), v(cols[2]), v(cols[3]),
                                                             v(d_o), v(d_c), 104 * n, sp, None)))
    N.check(lib.hlsmap_view_to_json_dev(ctx._h, n, v(cols[0]), v(cols[1]), v(cols[2]), v(cols[3]), v(d_o), v(d_c),
                                        104 * n, sp, ctypes.byref(total)))
    back = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(3)]
    b_t = torch.empty(n, dtype=torch.float64, device=dev)
    b_f, b_s = (torch.empty(n, dtype=torch.uint8, device=dev) for _ in range(2))
    ms_r = timed(lambda: N.check(lib.hlsmap_view_from_json_dev(ctx._h, n, v(d_c), v(d_o), v(back[0]), v(back[1]),
                                                               v(back[2]), v(b_t), v(b_f), v(b_s), sp)))
    exact = bool(torch.equal(back[2], cols[2]) and torch.equal(b_t, cols[3]) and int(b_s.sum()) == 0)
    alg = 24 * n + total.value
    out["views_json"] = {"views": n, "chars_per_view": total.value / n, "stringify_ms": ms_w,
                         "stringify_views_per_s": n / ms_w * 1e3, "stringify_roofline_frac": alg / ms_w / 1e6 / peak,
                         "parse_ms": ms_r, "parse_views_per_s": n / ms_r * 1e3,
                         "parse_roofline_frac": (26 * n + total.value) / ms_r / 1e6 / peak,   # text + 4 B offset in, 22 B of columns out
                         "round_trip_bit_exact": exact}
    return out



def list_form_sets(pk, synth, torch, dmap, dev, q0, rank, min_bytes):
    """Rotating query/result buffer sets for the list form: the same queries in other orders, so
    successive steps stream through distinct memory whose total is >= min_bytes (> 3x L2)."""
    nq = len(q0)
    sets = []
    while True:
        q = q0
        if sets:
            rng = synth._stream(7000 + rank, len(sets))
            perm = np.argsort(rng.u32(nq), kind="stable")
            q = pk.QueryBatch(q0.playlist[perm], q0.level[perm], q0.begin[perm], q0.duration[perm])
        _, total = dmap.count_segments(q.playlist, q.level, q.begin, q.duration)
        cap = int(total)
        sets.append({"q": q, "hits": cap,
                     "pl": torch.from_numpy(q.playlist).to(dev), "lv": torch.from_numpy(q.level).to(dev),
                     "bg": torch.from_numpy(q.begin).to(dev), "du": torch.from_numpy(q.duration).to(dev),
                     "off": torch.empty(nq + 1, dtype=torch.int32, device=dev),
                     "sn": torch.empty(max(cap, 1), dtype=torch.int32, device=dev),
                     "tm": torch.empty(max(cap, 1), dtype=torch.float64, device=dev)})
        per_step = 28 * nq + 12 * cap
        if len(sets) * per_step >= min_bytes and len(sets) >= 3:
            return sets, per_step


def timed_list_form(torch, ctx, dmap, sets, stream, steps, warmup, barrier=None, count_only=False):
    """`steps` back-to-back device-resident list-form batches over rotating buffer sets, CUDA events
    on the launch stream.  The K batches are handed to the library with ONE call
    (hlsmap_get_segment_list_dev_batches), so the host language's per-call cost cannot pace the GPU;
    with batch overlap allowed the library resolves up to 32 of them per kernel launch (one ramp and
    one tail per launch instead of per batch).  -> (region ms, mean hits per step, host time of that
    call per batch in us, (host t0, host t1) of the timed region, kernel launches of the call)."""
    from hlsjs_p2p_wrapper_b200 import _native as N
    nq = len(sets[0]["q"])

    def batch_array(n):
        arr = (N.Batch * n)()
        for i in range(n):
            s, a = sets[i % len(sets)], arr[i]
            a.nq, a.d_playlist, a.d_level = nq, s["pl"].data_ptr(), s["lv"].data_ptr()
            a.d_begin, a.d_duration, a.d_offsets = s["bg"].data_ptr(), s["du"].data_ptr(), s["off"].data_ptr()
            a.d_sn_out = None if count_only else s["sn"].data_ptr()
            a.d_time_out = None if count_only else s["tm"].data_ptr()
            a.hit_capacity, a.d_status = (0 if count_only else s["hits"]), None
        return arr
    warm, timed = batch_array(warmup), batch_array(steps)
    hits = sum(sets[i % len(sets)]["hits"] for i in range(steps))
    if warmup:
        assert dmap.get_segment_list_dev_batches(warm, stream) == warmup
    torch.cuda.synchronize()
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_begin.record()          # (creates the underlying cudaEvent_t handles; re-recorded below)
    t_end.record()
    torch.cuda.synchronize()
    if barrier:
        barrier()
    h0 = time.perf_counter()
    # both events are recorded on the launch stream from inside the call: t_begin immediately before
    # the first of the K launches, t_end immediately after the last
    done = dmap.get_segment_list_dev_batches(timed, stream, event_begin=t_begin.cuda_event, event_end=t_end.cuda_event)
    issue_us = (time.perf_counter() - h0) / steps * 1e6
    assert done == steps
    if barrier:
        barrier()
    else:
        torch.cuda.synchronize()
    h1 = time.perf_counter()
    launches = int(dmap.last_result(stream).kernel_launches)     # of the whole call, counted by the library
    return t_begin.elapsed_time(t_end), hits / steps, issue_us, (h0, h1), launches


def path_rows(pk, synth, torch, ctx, dmap, dev, tables, q0, stream, peak, steps, warmup):
    """Two more device-timed lines for the hot path on the same map (rank 0, N = 1): the look-ahead
    query class on its own (what the reference's caller asks for: ~30 s windows,
    hlsjs-p2p-wrapper-private.js:88-89) in list form, and the range form on the headline queries."""
    import ctypes
    from hlsjs_p2p_wrapper_b200 import _native as N
    out = {}
    nq = len(q0)
    table_bytes = 12 * len(tables.start)
    # -- look-ahead class: queries whose duration lies in the 29-31 s band of synth._queries
    _, big = synth.config2(seed=2, n_queries=int(nq * 1.6))
    sel = np.nonzero((big.duration >= 29.0) & (big.duration <= 31.0))[0][:nq]
    ql = pk.QueryBatch(big.playlist[sel], big.level[sel], big.begin[sel], big.duration[sel])
    if len(ql) == nq:
        sets, per_step = list_form_sets(pk, synth, torch, dmap, dev, ql, 0, 3 * L2_BYTES)
        ms, hits = timed_list_form(torch, ctx, dmap, sets, stream, steps, warmup)[:2]
        alg = 28.0 * nq + 12.0 * hits + table_bytes
        out["list_form_lookahead_mix"] = {
            "queries": nq, "hits_per_query": hits / nq, "ms_per_step": ms / steps,
            "queries_per_s": nq * steps / (ms * 1e-3), "algorithmic_bytes_per_launch": alg,
            "roofline_frac": alg / (ms / steps * 1e-3) / 1e9 / peak, "buffer_sets": len(sets),
            "what": "the 29-31 s duration class of the same generator (70 % of the headline mix), "
                    "list form, timed like the headline"}
        del sets
    # -- range form (first index + count per query), device-resident, rotating buffers: one batch
    #    size like the headline's and one large enough that launch overhead stops mattering
    lib = N.load()
    v = lambda x: ctypes.c_void_p(x.data_ptr())
    sp = ctypes.c_void_p(stream)
    for mult, key in ((1, "range_form_device"), (16, "range_form_device_16x")):
        n = nq * mult
        n_sets = max(2, int(np.ceil(3 * L2_BYTES / (33.0 * n))))
        rs = []
        for k in range(n_sets):
            idx = np.concatenate([np.argsort(synth._stream(9000 + k, j).u32(nq), kind="stable") for j in range(mult)])
            rs.append({"pl": torch.from_numpy(q0.playlist[idx]).to(dev), "lv": torch.from_numpy(q0.level[idx]).to(dev),
                       "bg": torch.from_numpy(q0.begin[idx]).to(dev), "du": torch.from_numpy(q0.duration[idx]).to(dev),
                       "first": torch.empty(n, dtype=torch.int32, device=dev),
                       "count": torch.empty(n, dtype=torch.int32, device=dev),
                       "st": torch.empty(n, dtype=torch.uint8, device=dev)})

        def launch(r):
            N.check(lib.hlsmap_get_segment_range_dev(dmap._h, n, v(r["pl"]), v(r["lv"]), v(r["bg"]), v(r["du"]),
                                                     v(r["first"]), v(r["count"]), v(r["st"]), sp))
        for i in range(warmup):
            launch(rs[i % n_sets])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            launch(rs[i % n_sets])
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        alg = 32.0 * n + table_bytes          # Q * (24 + 8) + sum F * 12 (the status byte is not counted)
        out[key] = {
            "queries": n, "ms_per_step": ms / steps, "queries_per_s": n * steps / (ms * 1e-3),
            "algorithmic_bytes_per_launch": alg, "roofline_frac": alg / (ms / steps * 1e-3) / 1e9 / peak,
            "buffer_sets": n_sets, "api": "hlsmap_get_segment_range_dev", "kernel": "range_small_kernel"}
        del rs
    return out


def sharded_config5(pk, synth, torch, dist, ctx, dev, rank, world, local, peak, n_chunks, gloo):
    """BASELINE.json configs[4]: 10^8 getSegmentList queries over 10^6 playlists, sharded by
    contiguous playlist range over the ranks (media-map.js:28,41: a query touches one playlist's one
    level).  Every rank plans the same boundaries from per-playlist counts (shard.plan_counts),
    generates only the corpus chunks its range touches, routes each chunk's queries by playlist
    (shard.split_queries), resolves them on its GPU with inputs resident (CUDA events around one
    launch per piece after a warm-up launch), and copies its results to the host; rank 0 gathers
    the shards and merges them (shard.merge) into the global list form.  No collective on the data
    path: the gather is host-side (gloo point-to-point of host arrays, or nothing at N = 1)."""
    import oracle
    from hlsjs_p2p_wrapper_b200 import shard
    P, QPP = synth.CONFIG5_CHUNK, synth.CONFIG5_QUERIES_PER_PLAYLIST
    t_all = time.perf_counter()
    frag_pp = np.concatenate([synth.config5_chunk_fragments_per_playlist(c) for c in range(n_chunks)])
    bounds = shard.plan_counts(frag_pp, np.full(len(frag_pp), QPP), world)
    p0, p1 = int(bounds[rank]), int(bounds[rank + 1])
    pieces = [c for c in range(n_chunks) if max(p0, c * P) < min(p1, (c + 1) * P)]
    stream = torch.cuda.Stream(device=dev)

    # generate the next chunk on a host thread while the GPU works on this one
    box = {}

    def gen(c):
        box[c] = synth.config5_chunk(c)
    th = None
    if pieces:
        th = threading.Thread(target=gen, args=(pieces[0],))
        th.start()
    dev_ms, d2h_s, gen_wait_s = 0.0, 0.0, 0.0
    n_q = n_hits = n_frag = 0
    alg = 0.0
    res_off, res_st, res_sn, res_tm = [], [], [], []
    sample = None
    ctx.set_batch_overlap(False)
    for k, c in enumerate(pieces):
        w0 = time.perf_counter()
        th.join()
        gen_wait_s += time.perf_counter() - w0
        t, q = box.pop(c)
        if k + 1 < len(pieces):
            th = threading.Thread(target=gen, args=(pieces[k + 1],))
            th.start()
        # this chunk's share of every rank's range; keep ours
        lb = np.clip(bounds - c * P, 0, P)
        qs, order = shard.split_queries(P, q, lb)
        a, b = int(lb[rank]), int(lb[rank + 1])
        tp, qp = t.slice_playlists(a, b), qs[rank]
        m = ctx.upload(tp)
        nq = len(qp)
        d = {kk: torch.from_numpy(getattr(qp, kk)).to(dev) for kk in ("playlist", "level", "begin", "duration")}
        _, total = m.count_segments(qp.playlist, qp.level, qp.begin, qp.duration)
        off = torch.empty(nq + 1, dtype=torch.int32, device=dev)
        st = torch.empty(nq, dtype=torch.uint8, device=dev)
        sn = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
        tm = torch.empty(max(total, 1), dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            for i in range(2):
                if i == 1:
                    e0.record()
                m.get_segment_list_dev(nq, d["playlist"].data_ptr(), d["level"].data_ptr(),
                                       d["begin"].data_ptr(), d["duration"].data_ptr(), off.data_ptr(),
                                       sn.data_ptr(), tm.data_ptr(), total, d_status=st.data_ptr(),
                                       stream=stream.cuda_stream)
            e1.record()
        torch.cuda.synchronize()
        dev_ms += e0.elapsed_time(e1)
        w0 = time.perf_counter()
        res_off.append(off.cpu().numpy())
        res_st.append(st.cpu().numpy())
        res_sn.append(sn[:total].cpu().numpy())
        res_tm.append(tm[:total].cpu().numpy())
        d2h_s += time.perf_counter() - w0
        if sample is None and rank == 0:
            # parity spot check against the CPU oracle on the first playlists of the first piece
            ns = min(2000, b - a)
            tsmp = tp.slice_playlists(0, ns)
            o_off, o_st, o_sn, o_tm = oracle.get_segment_list_batch(
                tsmp.playlist_base, tsmp.track_off, tsmp.track_state, tsmp.start, tsmp.sn,
                qp.playlist[:ns * QPP], qp.level[:ns * QPP], qp.begin[:ns * QPP], qp.duration[:ns * QPP])
            h = int(o_off[-1])
            sample = bool(np.array_equal(res_off[-1][:ns * QPP + 1].astype(np.int64), o_off) and
                          np.array_equal(res_sn[-1][:h], o_sn) and
                          np.array_equal(res_tm[-1][:h].view(np.uint64), o_tm.view(np.uint64)))
        n_q += nq
        n_hits += int(total)
        n_frag += tp.n_fragments
        alg += 28.0 * nq + 12.0 * total + 12.0 * tp.n_fragments
        m.close()
        del d, off, st, sn, tm
    # this rank's shard in list form: pieces are in playlist order already
    counts = np.concatenate([np.diff(o.astype(np.int64)) for o in res_off]) if res_off else np.zeros(0, np.int64)
    mine = (np.concatenate([[0], np.cumsum(counts)]),
            np.concatenate(res_st) if res_st else np.zeros(0, np.uint8),
            np.concatenate(res_sn) if res_sn else np.zeros(0, np.int32),
            np.concatenate(res_tm) if res_tm else np.zeros(0, np.float64))
    del res_off, res_st, res_sn, res_tm

    def gather_all(x):
        if world == 1:
            return [x]
        tt = torch.tensor([x], dtype=torch.float64, device=dev)
        outl = [torch.zeros_like(tt) for _ in range(world)]
        dist.all_gather(outl, tt)
        return [float(o.item()) for o in outl]
    per_rank_ms = gather_all(dev_ms)
    per_rank_q = gather_all(float(n_q))
    per_rank_hits = gather_all(float(n_hits))
    per_rank_alg = gather_all(alg)
    per_rank_frag = gather_all(float(n_frag))
    per_rank_d2h = gather_all(d2h_s)
    per_rank_genwait = gather_all(gen_wait_s)

    # host-side gather to rank 0 (host arrays over gloo), then shard.merge
    if world > 1:
        dist.barrier(group=gloo)
    g0 = time.perf_counter()
    merged = None
    if world == 1:
        shards = [mine]
    elif rank == 0:
        shards = [mine]
        for r in range(1, world):
            nqr, nhr = int(per_rank_q[r]), int(per_rank_hits[r])
            bufs = (torch.empty(nqr + 1, dtype=torch.int64), torch.empty(nqr, dtype=torch.uint8),
                    torch.empty(nhr, dtype=torch.int32), torch.empty(nhr, dtype=torch.float64))
            for x in bufs:
                dist.recv(x, src=r, group=gloo)
            shards.append(tuple(x.numpy() for x in bufs))
    else:
        for x in mine:
            dist.send(torch.from_numpy(np.ascontiguousarray(x)), dst=0, group=gloo)
        shards = None
    gather_s = time.perf_counter() - g0
    merge_s, ok_total = 0.0, None
    if rank == 0:
        g1 = time.perf_counter()
        total_q = int(sum(per_rank_q))
        merged = shard.merge(shards, np.arange(total_q))       # queries are grouped by playlist: shard order is query order
        merge_s = time.perf_counter() - g1
        ok_total = bool(int(merged[0][-1]) == int(sum(per_rank_hits)) and len(merged[2]) == int(sum(per_rank_hits))
                        and len(merged[0]) == total_q + 1)
        del merged, shards
    if world > 1:
        dist.barrier(group=gloo)
    if rank != 0:
        return None
    worst = max(per_rank_ms)
    total_q, total_alg = sum(per_rank_q), sum(per_rank_alg)
    return {
        "workload": "BASELINE.json configs[4]: %d getSegmentList queries over %d synthetic playlists (3 levels, "
                    "6-1024 fragments each, 100 queries per playlist, grouped by playlist), sharded by contiguous "
                    "playlist range over %d GPU(s); corpus = %d chunks of %d playlists, seeds %d.."
                    % (int(total_q), n_chunks * P, world, n_chunks, P, synth.CONFIG5_SEED0),
        "metric": METRIC, "unit": UNIT, "value": total_q / (worst * 1e-3), "scaling": "strong",
        "queries": int(total_q), "hits": int(sum(per_rank_hits)), "fragments": int(sum(per_rank_frag)),
        "hits_per_query": sum(per_rank_hits) / max(total_q, 1),
        "bounds": [int(x) for x in bounds],
        "per_rank_device_ms": per_rank_ms, "per_rank_queries": [int(x) for x in per_rank_q],
        "per_rank_algorithmic_bytes": per_rank_alg,
        "byte_imbalance_max_over_mean": max(per_rank_alg) / (total_alg / world),
        "algorithmic_bytes_per_query": total_alg / max(total_q, 1),
        "roofline_frac": total_alg / (worst * 1e-3) / 1e9 / (peak * world),
        "per_rank_roofline_frac": [a / (ms * 1e-3) / 1e9 / peak if ms > 0 else None
                                   for a, ms in zip(per_rank_alg, per_rank_ms)],
        "timing": "per piece (a chunk's share of the rank's playlist range, its own uploaded map): one warm-up "
                  "launch, then CUDA events around one launch on its stream, inputs and outputs resident; "
                  "summed per rank, value = all queries / max over ranks",
        "kernel": "resolve_search_kernel<SMALL=0> (tables of a piece: ~0.5 GB, staged per tile)",
        "d2h_s_max_over_ranks": max(per_rank_d2h), "host_gather_s": gather_s, "host_merge_s": merge_s,
        "gather": "none (one shard)" if world == 1 else "rank 0 receives every shard's host arrays over gloo "
                  "(point-to-point), then shard.merge",
        "merged_totals_ok": ok_total, "parity_sample_vs_oracle_ok": sample,
        "generation_wait_s_max_over_ranks": max(per_rank_genwait),
        "wall_s": time.perf_counter() - t_all,
    }


def run_ours(args):
    import torch
    import torch.distributed as dist
    import hlsjs_p2p_wrapper_b200 as pk
    from hlsjs_p2p_wrapper_b200 import synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: this engine has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    gloo = None
    # (diagnostics: HLSMAP_BENCH_PG1=1 brings NCCL up in a one-rank run, to see what its presence costs)
    pg1 = world == 1 and os.environ.get("HLSMAP_BENCH_PG1") == "1"
    if world > 1 or pg1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if pg1:
            os.environ.setdefault("MASTER_PORT", "29533")
            dist.init_process_group("nccl", device_id=dev, rank=0, world_size=1)
        else:
            dist.init_process_group("nccl", device_id=dev)
            gloo = dist.new_group(backend="gloo")      # host-side gather of the sharded leg only

    def barrier():
        if world > 1 or pg1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def all_ranks(x):
        if world == 1:
            return [x]
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        out = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [float(o.item()) for o in out]

    # clocks: the sampler needs ~0.5 s to bring NVML up; start it before anything else so that
    # even a 1 ms timed region has samples on either side of it and, with luck, inside
    try:
        uuid = str(torch.cuda.get_device_properties(local).uuid)
    except Exception:
        uuid = None
    sampler = ClockSampler(uuid, local)
    if rank == 0:
        sampler.start()

    ctx = pk.Context(local)
    # every rank owns its own playlist (seeded by rank): sharding by playlist, no exchange
    tables, q0 = synth.config2(seed=2 + 1000 * rank, n_queries=args.queries)
    dmap = ctx.upload(tables)
    info = dmap.info()
    nq = args.queries

    # Rotating query/result buffer sets: successive steps stream through distinct memory
    # whose total is > 3x L2, so no step finds its inputs or outputs in L2.
    sets, per_step_bytes = list_form_sets(pk, synth, torch, dmap, dev, q0, rank, 3 * L2_BYTES)
    n_sets = len(sets)
    torch.cuda.synchronize()
    # an explicit stream: the launches and the CUDA events that time them share it
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0

    count_only = os.environ.get("HLSMAP_BENCH_COUNT_ONLY") == "1"     # experiment switch

    s0 = sets[0]
    dmap.get_segment_list_dev(nq, s0["pl"].data_ptr(), s0["lv"].data_ptr(), s0["bg"].data_ptr(),
                              s0["du"].data_ptr(), s0["off"].data_ptr(), s0["sn"].data_ptr(),
                              s0["tm"].data_ptr(), s0["hits"], stream=stream)
    chk = dmap.last_result(stream)
    assert chk.total_hits == s0["hits"]

    if rank == 0:
        sampler.wait_first(10.0)
    # Consecutive steps use disjoint buffers, so the engine may start step i+1 while step i is
    # still writing its last hits (programmatic dependent launch): a batch's ramp-up then hides
    # behind its predecessor's tail, as it would in a serving loop.
    overlap = os.environ.get("HLSMAP_BENCH_NO_OVERLAP") != "1"
    ctx.set_batch_overlap(overlap)
    my_ms, hits_per_step, issue_us, sampler.window, timed_launches = timed_list_form(
        torch, ctx, dmap, sets, stream, args.steps, args.warmup, barrier=barrier, count_only=count_only)
    region_ms = max_over_ranks(my_ms)
    per_rank_ms = all_ranks(my_ms / args.steps)
    per_rank_issue_us = all_ranks(issue_us)
    total_queries = sum_over_ranks(float(nq * args.steps))
    value = total_queries / (region_ms * 1e-3)
    mean_kernel_ms = my_ms / args.steps     # average launch-to-launch time on this rank

    # the same batches as one launch per batch (overlapping launches, no fusing), for reference
    unfused_ms = None
    if overlap and not count_only:
        ctx.set_batch_overlap(2)
        unfused_ms = timed_list_form(torch, ctx, dmap, sets, stream, args.steps, args.warmup)[0] / args.steps
        ctx.set_batch_overlap(overlap)

    # the same kernel timed one launch at a time (no overlap), for reference
    ctx.set_batch_overlap(False)
    iso_steps = min(args.steps, 200)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(iso_steps)]
    torch.cuda.synchronize()
    for i in range(iso_steps):
        s = sets[i % n_sets]
        ev[i][0].record()
        dmap.get_segment_list_dev(nq, s["pl"].data_ptr(), s["lv"].data_ptr(), s["bg"].data_ptr(),
                                  s["du"].data_ptr(), s["off"].data_ptr(), s["sn"].data_ptr(),
                                  s["tm"].data_ptr(), s["hits"], stream=stream)
        ev[i][1].record()
    torch.cuda.synchronize()
    kernel_ms = [a.elapsed_time(b) for a, b in ev]
    ctx.set_batch_overlap(overlap)
    clocks = sampler.stop() if rank == 0 else None

    # roofline of the dominant (only) kernel: algorithmic bytes / mean launch duration
    table_bytes = 12 * info["n_fragments"]
    alg_bytes = 28.0 * nq + 12.0 * hits_per_step + table_bytes
    achieved = alg_bytes / (mean_kernel_ms * 1e-3) / 1e9
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("resolve_search_kernel_dram_bytes_per_launch")
        except Exception:
            traffic = None

    # end to end through the host-pointer C ABI call on pinned buffers
    pins = {k: ctx.pinned(shape, dt) for k, shape, dt in (
        ("pl", nq, np.int32), ("lv", nq, np.int32), ("bg", nq, np.float64), ("du", nq, np.float64),
        ("off", nq + 1, np.int32), ("sn", s0["hits"], np.int32), ("tm", s0["hits"], np.float64))}
    pins["pl"].array[:] = s0["q"].playlist
    pins["lv"].array[:] = s0["q"].level
    pins["bg"].array[:] = s0["q"].begin
    pins["du"].array[:] = s0["q"].duration

    def e2e_step():
        return dmap.get_segment_list_into(pins["pl"].array, pins["lv"].array, pins["bg"].array,
                                          pins["du"].array, pins["off"].array, pins["sn"].array,
                                          pins["tm"].array)
    for _ in range(3):
        r = e2e_step()
    e2e_launches = r.kernel_launches
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        r = e2e_step()
    torch.cuda.synchronize()
    my_e2e_s = time.perf_counter() - t0
    e2e_s = max_over_ranks(my_e2e_s)
    per_rank_e2e_ms = all_ranks(1e3 * my_e2e_s / args.e2e_steps)
    e2e_value = sum_over_ranks(float(nq * args.e2e_steps)) / e2e_s
    assert int(pins["off"].array[-1]) == s0["hits"] == r.total_hits
    h2d = 24 * nq * world                               # all ranks, per step
    d2h = (4 * (nq + 1) + 12 * s0["hits"]) * world

    # the same queries through the range form (first index + count per query: 9 B back instead of
    # 12 B per hit) -- what a PCIe-bound deployment would call; reported beside e2e, not instead of it
    import ctypes
    from hlsjs_p2p_wrapper_b200 import _native as N
    rp = {k: ctx.pinned(nq, dt) for k, dt in (("first", np.int32), ("count", np.int32), ("st", np.uint8))}
    pv = lambda a: ctypes.c_void_p(a.ctypes.data)

    def range_step():
        N.check(N.load().hlsmap_get_segment_range(dmap._h, nq, pv(pins["pl"].array), pv(pins["lv"].array),
                                                  pv(pins["bg"].array), pv(pins["du"].array),
                                                  pv(rp["first"].array), pv(rp["count"].array), pv(rp["st"].array)))
    for _ in range(3):
        range_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        range_step()
    dt_r = max_over_ranks(time.perf_counter() - t0)
    assert int(rp["count"].array.astype(np.int64).sum()) == s0["hits"]
    e2e_range = {"value": nq * args.e2e_steps * world / dt_r, "unit": UNIT, "ms_per_step": 1e3 * dt_r / args.e2e_steps,
                 "h2d_bytes_per_step": 24 * nq * world, "d2h_bytes_per_step": 9 * nq * world,
                 "api": "hlsmap_get_segment_range (host pointers, pinned)"}
    for v in rp.values():
        v.free()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = host_threads()
        rate, reps, dt, _ = cpu_filter_rate(tables, q0, threads, min_seconds=3.0)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": "%d x the full %d-query batch (%.1f s wall) through the C restatement of "
                         "media-map.js:42-51 with %d pthreads; node on this box: %s"
                         % (reps, nq, dt, threads, node_version() or "absent")}

    rows = None
    if rank == 0 and world == 1 and not args.no_other_rows:
        try:
            rows = path_rows(pk, synth, torch, ctx, dmap, dev, tables, q0, stream, peak,
                             min(args.steps, 200), max(args.warmup, 3))
            rows.update(other_rows(ctx, torch, peak))
        except Exception as e:                      # never let the extra rows cost the headline line
            rows = dict(rows or {}, error="%s: %s" % (type(e).__name__, e))

    for p in pins.values():
        p.free()
    del sets
    sharded = None
    if not args.no_sharded:
        try:
            sharded = sharded_config5(pk, synth, torch, dist, ctx, dev, rank, world, local, peak,
                                      args.config5_chunks, gloo)
        except Exception as e:
            if world > 1:
                raise
            sharded = {"error": "%s: %s" % (type(e).__name__, e)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": region_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, world),
            "details": {"l2_policy": "rotating %d distinct query/result buffer sets, %.0f MB per step, "
                                     "%.0f MB per rotation (> 3x the 126 MB L2): no step finds its "
                                     "inputs or outputs in L2" % (n_sets, per_step_bytes / 1e6,
                                                                  n_sets * per_step_bytes / 1e6),
                        "hits_per_query": hits_per_step / nq, "kernel": "resolve_search_kernel",
                        "stage_all": info["stage_all"],
                        "per_rank_ms_per_step": per_rank_ms,
                        "host_issue_us_per_launch": per_rank_issue_us,
                        "host_issue_note": "time.perf_counter around the timed launch loop / steps, per rank: the "
                                           "Python -> ctypes -> cudaLaunchKernelEx path; the GPU is the limiter "
                                           "while this stays below ms_per_step"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "frac_of_nominal_8TBs": achieved / 8000.0,
                         "algorithmic_bytes_per_batch": alg_bytes,
                         "algorithmic_bytes_per_launch": alg_bytes * args.steps / max(timed_launches, 1),
                         "kernel_ms_mean": mean_kernel_ms,
                         "timing": "CUDA events on the launch stream around the %d timed batches, recorded from "
                                   "inside the one C call that issues them (begin: right before the first launch, "
                                   "end: right after the last); consecutive batches may overlap (the caller "
                                   "promised disjoint buffers) = %s: the library then resolves up to 32 batches "
                                   "per kernel launch (%d launches for the %d batches) and lets consecutive "
                                   "launches overlap (programmatic dependent launch); kernel_ms_mean = region / "
                                   "batches" % (args.steps, overlap, timed_launches, args.steps),
                         "kernel_launches_timed_region": timed_launches,
                         "one_launch_per_batch": None if unfused_ms is None else {
                             "kernel_ms_mean": unfused_ms, "frac": alg_bytes / (unfused_ms * 1e-3) / 1e9 / peak,
                             "note": "the same %d batches through the same call with "
                                     "hlsmap_ctx_set_batch_overlap(ctx, 2): consecutive launches overlap, "
                                     "one launch per batch (rank 0)" % args.steps},
                         "kernel_ms_isolated_mean": float(np.mean(kernel_ms)),
                         "kernel_ms_isolated_min": float(np.min(kernel_ms)),
                         "frac_isolated": alg_bytes / (float(np.mean(kernel_ms)) * 1e-3) / 1e9 / peak,
                         "isolated_note": "one launch at a time with the overlap option off (same kernel, same "
                                          "512-query tiles): a lone batch pays its own ramp-up and tail"},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": args.e2e_steps,
                    "ms_per_step": 1e3 * e2e_s / args.e2e_steps,
                    "per_rank_ms_per_step": per_rank_e2e_ms,
                    "api": "hlsmap_get_segment_list (host pointers, pinned)",
                    "kernel_launches_per_step": e2e_launches,
                    "capacity_note": "the caller's hit buffers were sized from an earlier, untimed count pass "
                                     "(hlsmap_get_segment_list with NULL hit arrays); a caller that cannot know "
                                     "the count either pays that pass (24 B/query up, 4 B/query back) or "
                                     "over-allocates"},
            "gpu_launches": timed_launches * world, "clocks": clocks, "node": node_version(),
            "other_rows": rows, "e2e_range_form": e2e_range, "sharded_config5": sharded,
        }
        print(json.dumps(line))
    dmap.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

"""Forty mixed-size list-form batches through hlsmap_get_segment_list_dev_batches, alternately fused (overlap mode 1)
and as one overlapping launch per batch (mode 2); prints which batches differ from the oracle and which were left
untouched.  (How the out-of-order hand-over of the scratch sets was found; tests/test_gpu_parity.py now covers it.)"""
import sys, os, ctypes
import numpy as np, torch
sys.path.insert(0, "/root/repo")
import oracle
import hlsjs_p2p_wrapper_b200 as pk
from hlsjs_p2p_wrapper_b200 import synth, _native as N
from hlsjs_p2p_wrapper_b200.tables import QueryBatch
dev = torch.device("cuda:0")
ctx = pk.Context(0)
t, _ = synth.config2(n_queries=4)
m = ctx.upload(t)
stream = torch.cuda.Stream(device=dev)
sizes = [30_000, 513, 1, 70_001, 0, 1024, 2_049, 120_000] * 5
sets = []
for i, nq in enumerate(sizes):
    q = synth.config2(seed=300 + i, n_queries=max(nq, 1))[1]
    q = QueryBatch(q.playlist[:nq], q.level[:nq], q.begin[:nq], q.duration[:nq])
    ref = oracle.get_segment_list_batch(t.playlist_base, t.track_off, t.track_state, t.start, t.sn, q.playlist, q.level, q.begin, q.duration, threads=8)
    H = int(ref[0][-1])
    sets.append({"ref": ref, "H": H, "nq": nq,
                 "pl": torch.from_numpy(q.playlist).to(dev), "lv": torch.from_numpy(q.level).to(dev),
                 "bg": torch.from_numpy(q.begin).to(dev), "du": torch.from_numpy(q.duration).to(dev),
                 "off": torch.full((nq + 1,), -1, dtype=torch.int32, device=dev),
                 "sn": torch.full((max(H, 1) + 256,), -7, dtype=torch.int32, device=dev),
                 "tm": torch.full((max(H, 1) + 256,), -7.0, dtype=torch.float64, device=dev)})
def batch(s):
    return {"nq": s["nq"], "d_playlist": s["pl"].data_ptr(), "d_level": s["lv"].data_ptr(),
            "d_begin": s["bg"].data_ptr(), "d_duration": s["du"].data_ptr(), "d_offsets": s["off"].data_ptr(),
            "d_sn_out": s["sn"].data_ptr(), "d_time_out": s["tm"].data_ptr(), "capacity": s["H"]}
for mode in (1, 1, 2, 2, 1, 2, 2):
    ctx.set_batch_overlap(mode)
    for s in sets:
        s["off"].fill_(-1)
    torch.cuda.synchronize()
    n = m.get_segment_list_dev_batches([batch(s) for s in sets], stream.cuda_stream)
    last = m.last_result(stream.cuda_stream)
    torch.cuda.synchronize()
    bad = [i for i, s in enumerate(sets) if not np.array_equal(s["off"].cpu().numpy().astype(np.int64), s["ref"][0])]
    untouched = [i for i in bad if int(sets[i]["off"][0].item()) == -1]
    print("mode", mode, "issued", n, "launches", last.kernel_launches, "bad", bad, "untouched", untouched, flush=True)

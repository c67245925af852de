"""Sharding by playlist across GPUs: a query touches exactly one
playlist's one level (media-map.js:28,41), playlists share nothing, so each GPU holds a
contiguous playlist range and the only cross-GPU step is a host-side gather.  No NCCL.

`plan` / `split` / `merge` are pure host logic (tested on CPU with gloo, world_size 2);
`ShardedMap` drives one Context per GPU from one process with a thread per GPU.
"""
import threading

import numpy as np

from .tables import FragmentTables, QueryBatch


def plan_counts(frag_per_pl, q_per_pl, n_shards: int):
    """`plan` from per-playlist counts alone (fragments, queries): the tables themselves are not
    needed to place the boundaries, so a rank can plan before it has generated or loaded anything."""
    frag_per_pl = np.asarray(frag_per_pl, np.float64)
    p = len(frag_per_pl)
    w = 12.0 * frag_per_pl + 76.0 * np.asarray(q_per_pl, np.float64) + 1e-9
    c = np.concatenate([[0.0], np.cumsum(w)])
    targets = c[-1] * np.arange(1, n_shards) / n_shards
    cuts = np.searchsorted(c, targets, side="left")
    bounds = np.concatenate([[0], cuts, [p]]).astype(np.int64)
    return np.maximum.accumulate(bounds)


def plan(tables: FragmentTables, queries: QueryBatch, n_shards: int):
    """Contiguous playlist ranges balanced by bytes moved: 12 B per fragment of the shard's
    tables + 76 B per query.  Returns playlist boundaries, length n_shards+1."""
    p = tables.n_playlists
    frag_per_pl = np.diff(tables.track_off[tables.playlist_base].astype(np.int64))
    inb = (queries.playlist >= 0) & (queries.playlist < p)
    q_per_pl = np.bincount(queries.playlist[inb], minlength=p)[:p]
    return plan_counts(frag_per_pl, q_per_pl, n_shards)


def split_queries(n_playlists: int, queries: QueryBatch, bounds):
    """-> (query_shards, order): query_shards[i] = the queries of shard i with playlist ids re-based;
    `order` is the permutation that groups the original queries shard by shard.  Queries on
    playlists outside [0, P) go to shard 0 (they resolve to 'missing' there too)."""
    n = len(bounds) - 1
    pl = queries.playlist.astype(np.int64)
    shard_of = np.clip(np.searchsorted(bounds, pl, side="right") - 1, 0, n - 1)
    outside = (pl < 0) | (pl >= n_playlists)
    shard_of[outside] = 0
    order = np.argsort(shard_of, kind="stable")
    counts = np.bincount(shard_of, minlength=n)
    starts = np.concatenate([[0], np.cumsum(counts)])
    out = []
    for i in range(n):
        idx = order[starts[i]:starts[i + 1]]
        lp = pl[idx] - int(bounds[i])
        if i == 0:
            lp = np.where(outside[idx], -1, lp)
        out.append(QueryBatch(lp.astype(np.int32), queries.level[idx].copy(), queries.begin[idx].copy(),
                              queries.duration[idx].copy()))
    return out, order


def split(tables: FragmentTables, queries: QueryBatch, bounds):
    """-> (shards, order): shards[i] = (tables_i, queries_i), both re-based to the shard's first
    playlist; `order` as in split_queries.  (Slices the tables: call it once per upload, and
    split_queries per batch.)"""
    qs, order = split_queries(tables.n_playlists, queries, bounds)
    shards = [(tables.slice_playlists(int(bounds[i]), int(bounds[i + 1])), q) for i, q in enumerate(qs)]
    return shards, order


def split_texts(playlist_base, track_state, text, text_off, bounds):
    """Playlists still in text form, cut on the same playlist boundaries: shard i gets
    (playlist_base_i int32, track_state_i uint8, text_i uint8 view, text_off_i int64), ready for
    `Context.upload_m3u8`.  The text of a shard is one contiguous slice: nothing is copied."""
    playlist_base = np.asarray(playlist_base, np.int64)
    text_off = np.asarray(text_off, np.int64)
    text = np.frombuffer(text, np.uint8) if isinstance(text, (bytes, bytearray, memoryview)) else np.asarray(text, np.uint8)
    out = []
    for i in range(len(bounds) - 1):
        p0, p1 = int(bounds[i]), int(bounds[i + 1])
        t0, t1 = int(playlist_base[p0]), int(playlist_base[p1])
        b0, b1 = int(text_off[t0]), int(text_off[t1])
        out.append(((playlist_base[p0:p1 + 1] - t0).astype(np.int32), np.asarray(track_state[t0:t1], np.uint8),
                    text[b0:b1], text_off[t0:t1 + 1] - b0))
    return out


def merge(results, order):
    """Host-side gather: per-shard list-form results (offsets, status, sn, time) back into
    the original query order."""
    counts = np.concatenate([np.diff(r[0]).astype(np.int64) for r in results])
    status = np.concatenate([r[1] for r in results])
    sn = np.concatenate([r[2] for r in results])
    tm = np.concatenate([r[3] for r in results])
    nq = len(order)
    if nq == 0 or (order[0] == 0 and order[-1] == nq - 1 and np.all(np.diff(order) == 1)):
        # queries arrived grouped by shard (e.g. by playlist): shard order IS query order
        return np.concatenate([[0], np.cumsum(counts)]), status, sn, tm
    inv = np.empty(nq, np.int64)
    inv[order] = np.arange(nq)
    src_off = np.concatenate([[0], np.cumsum(counts)])
    out_counts = counts[inv]
    out_off = np.concatenate([[0], np.cumsum(out_counts)])
    # gather each query's run of hits
    take = np.repeat(src_off[:-1][inv] - out_off[:-1], out_counts) + np.arange(out_off[-1])
    return out_off, status[inv], sn[take], tm[take]


class ShardedMap:
    """Tables of many playlists spread over several GPUs of one box, one Context each."""

    def __init__(self, contexts, tables: FragmentTables, queries_hint: QueryBatch = None):
        self.contexts = list(contexts)
        hint = queries_hint or QueryBatch(np.zeros(0, np.int32), np.zeros(0, np.int32),
                                          np.zeros(0), np.zeros(0))
        self.bounds = plan(tables, hint, len(self.contexts))
        self.tables = tables
        self.maps = [c.upload(tables.slice_playlists(int(self.bounds[i]), int(self.bounds[i + 1])))
                     for i, c in enumerate(self.contexts)]

    @classmethod
    def from_m3u8(cls, contexts, playlist_base, track_state, text, text_off):
        """Each GPU parses its own playlists' text; bounds split the text evenly by bytes (queries
        are routed by playlist id alone, so the host never needs the parsed tables)."""
        self = object.__new__(cls)
        self.contexts = list(contexts)
        pb, to = np.asarray(playlist_base, np.int64), np.asarray(text_off, np.int64)
        c = to[pb].astype(np.float64)                       # text bytes before each playlist
        cuts = np.searchsorted(c, c[-1] * np.arange(1, len(self.contexts)) / len(self.contexts), side="left")
        self.bounds = np.maximum.accumulate(np.concatenate([[0], cuts, [len(pb) - 1]]).astype(np.int64))
        nt = len(to) - 1
        self.tables = FragmentTables(pb.astype(np.int32), np.zeros(nt + 1, np.int64), np.asarray(track_state, np.uint8),
                                     np.ones(nt, np.int32), np.zeros(0), np.zeros(0, np.int32), np.zeros(0))
        parts = split_texts(pb, track_state, text, to, self.bounds)
        self.maps = [c.upload_m3u8(b, st, tx, off) for c, (b, st, tx, off) in zip(self.contexts, parts)]
        return self

    def get_segment_list(self, queries: QueryBatch):
        """One batch over all GPUs: route by playlist, one host thread per GPU, gather on the host.
        An error on any GPU (capacity, out of memory, ...) is re-raised here."""
        qs, order = split_queries(self.tables.n_playlists, queries, self.bounds)
        out = [None] * len(self.maps)
        errors = [None] * len(self.maps)

        def work(i):
            try:
                q = qs[i]
                r = self.maps[i].get_segment_list(q.playlist, q.level, q.begin, q.duration)
                out[i] = (r.offsets.astype(np.int64), r.status, r.sn, r.time)
            except BaseException as e:      # handed to the caller below
                errors[i] = e
        threads = [threading.Thread(target=work, args=(i,)) for i in range(len(self.maps))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        for e in errors:
            if e is not None:
                raise e
        return merge(out, order)

    def close(self):
        for m in self.maps:
            m.close()

"""Host-side SoA form of `hls.levels[..].details.fragments` for many playlists.

The reference reads hls.js level objects directly (media-map.js:28,41,43-48,61,66-69,
82-84; fixture test/mocks/hls.js:12-45).  The batch engine needs the same information as
flat columns; this module converts between the two and slices tables by playlist range
for multi-GPU sharding.  Pure numpy: no arithmetic of the hot path happens here.
"""
from dataclasses import dataclass

import numpy as np

TRACK_MISSING, TRACK_UNPARSED, TRACK_PARSED = 0, 1, 2


def _field(obj, key, default=None):
    if obj is None:
        return default
    if isinstance(obj, dict):
        return obj.get(key, default)
    return getattr(obj, key, default)


@dataclass
class FragmentTables:
    playlist_base: np.ndarray   # int32[P+1]  first track of each playlist
    track_off: np.ndarray       # int64[T+1]  first fragment of each track
    track_state: np.ndarray     # uint8[T]    TRACK_*
    url_count: np.ndarray       # int32[T]    level.url.length
    start: np.ndarray           # float64[F]  fragment.start
    sn: np.ndarray              # int32[F]    fragment.sn
    duration: np.ndarray        # float64[F]  fragment.duration (NaN where absent)

    @property
    def n_playlists(self):
        return len(self.playlist_base) - 1

    @property
    def n_tracks(self):
        return len(self.track_off) - 1

    @property
    def n_fragments(self):
        return int(self.track_off[-1])

    def validate(self):
        assert self.playlist_base.dtype == np.int32 and self.track_off.dtype == np.int64
        assert self.playlist_base[0] == 0 and self.playlist_base[-1] == self.n_tracks
        assert self.track_off[0] == 0 and np.all(np.diff(self.track_off) >= 0)
        assert len(self.start) == len(self.sn) == len(self.duration) == self.n_fragments
        return self

    @staticmethod
    def from_hls_levels(playlists):
        """`playlists`: a list with one entry per player, each the value of `hls.levels`
        (a list of level objects/dicts with `url`, optional `details.fragments` of
        `{sn, start[, duration]}`) or None/undefined when the master playlist is not parsed.
        """
        pbase, toff, state, urlc = [0], [0], [], []
        start, sn, dur = [], [], []
        for levels in playlists:
            levels = levels if levels else []
            for level in levels:
                if not level:
                    state.append(TRACK_MISSING)
                    urlc.append(0)
                    toff.append(toff[-1])
                    continue
                url = _field(level, "url", [])
                urlc.append(len(url) if url is not None else 0)
                details = _field(level, "details")
                frags = _field(details, "fragments") if details else None
                if not details:
                    state.append(TRACK_UNPARSED)
                    toff.append(toff[-1])
                    continue
                state.append(TRACK_PARSED)
                frags = frags if frags is not None else []
                for f in frags:
                    start.append(float(_field(f, "start")))
                    sn.append(int(_field(f, "sn")))
                    d = _field(f, "duration")
                    dur.append(float("nan") if d is None else float(d))
                toff.append(toff[-1] + len(frags))
            pbase.append(len(state))
        return FragmentTables(
            np.asarray(pbase, np.int32), np.asarray(toff, np.int64), np.asarray(state, np.uint8),
            np.asarray(urlc, np.int32), np.asarray(start, np.float64), np.asarray(sn, np.int32),
            np.asarray(dur, np.float64)).validate()

    def slice_playlists(self, p0, p1):
        """Tables of playlists [p0, p1) re-based to start at zero (one GPU's shard)."""
        t0, t1 = int(self.playlist_base[p0]), int(self.playlist_base[p1])
        f0, f1 = int(self.track_off[t0]), int(self.track_off[t1])
        return FragmentTables(
            (self.playlist_base[p0:p1 + 1] - t0).astype(np.int32),
            (self.track_off[t0:t1 + 1] - f0).astype(np.int64),
            self.track_state[t0:t1].copy(), self.url_count[t0:t1].copy(),
            self.start[f0:f1].copy(), self.sn[f0:f1].copy(), self.duration[f0:f1].copy())

    def dur0(self):
        """First-fragment duration per track (what getSegmentDuration returns)."""
        out = np.full(self.n_tracks, np.nan)
        lens = np.diff(self.track_off)
        ok = (lens > 0) & (self.track_state == TRACK_PARSED)
        out[ok] = self.duration[self.track_off[:-1][ok]]
        return out


@dataclass
class QueryBatch:
    playlist: np.ndarray   # int32[Q]
    level: np.ndarray      # int32[Q]
    begin: np.ndarray      # float64[Q]
    duration: np.ndarray   # float64[Q]

    def __len__(self):
        return len(self.level)

    def slice(self, q0, q1, playlist_shift=0):
        return QueryBatch((self.playlist[q0:q1] - playlist_shift).astype(np.int32),
                          self.level[q0:q1].copy(), self.begin[q0:q1].copy(),
                          self.duration[q0:q1].copy())

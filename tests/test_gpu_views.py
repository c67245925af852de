"""SegmentView / TrackView batch operations on the GPU against the oracle: toArrayBuffer,
fromArrayBuffer, isEqual, isInTrack, the fused round trip (BASELINE.json configs[3]) and
viewToString.  Byte-exact."""
import numpy as np
import pytest

import oracle
from oracle import pyref
from hlsjs_p2p_wrapper_b200 import synth

pytestmark = pytest.mark.gpu


def _views(n, seed=0, negatives=False):
    rng = np.random.default_rng(seed)
    lo = -(1 << 31) if negatives else 0
    level = rng.integers(lo, 1 << 31, n).astype(np.int32)
    url = rng.integers(lo, 1 << 31, n).astype(np.int32)
    sn = rng.integers(-(1 << 31), 1 << 31, n).astype(np.int32)
    small = rng.random(n) < 0.5
    level[small] = rng.integers(0, 8, small.sum())
    url[small] = rng.integers(0, 3, small.sum())
    sn[small] = rng.integers(0, 12000, small.sum())
    return level, url, sn


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 1023, 1024, 1025, 100_003])
def test_pack_unpack_equal(ctx, n):
    level, url, sn = _views(n, seed=n)
    wire = ctx.view_pack(level, url, sn)
    assert wire.tobytes() == oracle.view_pack(level, url, sn).tobytes()
    lv, ui, s = ctx.view_unpack(wire)
    elv, eui, es = oracle.view_unpack(wire)
    assert np.array_equal(lv, elv) and np.array_equal(ui, eui) and np.array_equal(s, es)
    l2, u2, s2 = _views(n, seed=n + 1)
    mix = np.arange(n) % 3 == 0
    l2[mix], u2[mix], s2[mix] = level[mix], url[mix], sn[mix]
    assert np.array_equal(ctx.view_is_equal(level, url, sn, l2, u2, s2),
                          oracle.view_is_equal(level, url, sn, l2, u2, s2))
    assert np.array_equal(ctx.view_is_in_track(level, url, l2, u2),
                          oracle.view_is_in_track(level, url, l2, u2))


def test_first_view_matches_reference_bytes(ctx, golden):
    v = golden["segmentView"]["arraybuffer"][0]
    wire = ctx.view_pack([v["ctor"]["trackView"]["level"]], [v["ctor"]["trackView"]["urlId"]], [v["ctor"]["sn"]])
    assert wire.tobytes().hex() == v["expect_bytes_hex"]
    sv = pyref.SegmentView(v["ctor"])
    assert wire.tobytes() == sv.toArrayBuffer()


def test_config4_roundtrip(ctx):
    """10^6 views drawn as getSegmentList hits on 8 levels x 10k fragments: SoA -> 12-byte
    wire -> SoA -> isEqual, with 1 % perturbed so `false` occurs."""
    t, views, other = synth.config4()
    wire, eq = ctx.view_roundtrip(views["level"], views["urlId"], views["sn"])
    ewire, eeq = oracle.view_roundtrip(views["level"], views["urlId"], views["sn"])
    assert wire.tobytes() == ewire.tobytes() and np.array_equal(eq, eeq) and eq.all()
    lv, ui, sn = ctx.view_unpack(wire)
    same = ctx.view_is_equal(lv.astype(np.int32), ui.astype(np.int32), sn.astype(np.int32),
                             other["level"], other["urlId"], other["sn"])
    expect = oracle.view_is_equal(views["level"], views["urlId"], views["sn"],
                                  other["level"], other["urlId"], other["sn"])
    assert np.array_equal(same, expect)
    assert 0 < (same == 0).sum() < 0.02 * len(same)
    # the views really are hits of the uploaded tables
    m = ctx.upload(t)
    pick = np.arange(0, len(sn), 997)
    res = m.get_segment_list(np.zeros(len(pick), np.int32), views["level"][pick], views["time"][pick],
                             np.zeros(len(pick)))
    assert (np.diff(res.offsets) >= 1).all()
    assert np.array_equal(res.sn[res.offsets[:-1]], views["sn"][pick])
    m.close()


def test_roundtrip_negative_fields_do_not_survive(ctx):
    """ToUint32 wraps negatives, the parsed side reads them back as 2^32 - |v|: not equal."""
    level, url, sn = _views(50_001, seed=9, negatives=True)
    wire, eq = ctx.view_roundtrip(level, url, sn)
    ewire, eeq = oracle.view_roundtrip(level, url, sn)
    assert wire.tobytes() == ewire.tobytes() and np.array_equal(eq, eeq)
    assert (eq == 0).any() and (eq == 1).any()
    i = int(np.nonzero(sn < 0)[0][0])
    sv = pyref.SegmentView({"sn": int(sn[i]), "trackView": {"level": int(level[i]), "urlId": int(url[i])}})
    assert pyref.SegmentView.fromArrayBuffer(sv.toArrayBuffer()).isEqual(sv) is False and eq[i] == 0


@pytest.mark.parametrize("n", [0, 1, 5, 1024, 1025, 70_001])
@pytest.mark.parametrize("with_sn", [True, False])
def test_view_to_string(ctx, n, with_sn):
    level, url, sn = _views(n, seed=3 + n, negatives=True)
    if n >= 5:
        level[:5] = [0, -1, 2147483647, -2147483648, 10]
        sn[:5] = [0, 9, 10, -2147483648, 1000000000]
    off, chars = ctx.view_format(level, url, sn, with_sn=with_sn)
    eoff, echars = oracle.view_format(level, url, sn, with_sn=with_sn)
    assert np.array_equal(off.astype(np.int64), eoff) and chars.tobytes() == echars.tobytes()
    for i in range(min(n, 64)):
        sv = pyref.SegmentView({"sn": int(sn[i]), "trackView": {"level": int(level[i]), "urlId": int(url[i])}})
        want = sv.viewToString() if with_sn else sv.trackView.viewToString()
        assert chars[off[i]:off[i + 1]].tobytes().decode() == want

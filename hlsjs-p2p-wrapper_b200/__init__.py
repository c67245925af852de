"""B200-native batch engine for the MediaMap / SegmentView / TrackView path of
streamroot-hlsjs-p2p-wrapper (lib/integration/mapping/*.js).

Layout: ``csrc/`` holds the sm_100a CUDA kernels and the C ABI (``include/hlsmap.h``);
``engine`` is the numpy-facing binding; ``mapping`` mirrors the reference's JavaScript
classes (same names, arguments and errors) on top of it; ``synth`` generates the
BASELINE.json workloads; ``shard`` splits work by playlist across GPUs.
Importing the package does not need a GPU; creating a Context does.
"""
from . import _native
from .tables import FragmentTables, QueryBatch
from .engine import Context, DeviceMap, SegmentListResult
from .mapping import MediaMap, SegmentView, TrackView, JSError, JSTypeError, undefined

__all__ = ["Context", "DeviceMap", "FragmentTables", "QueryBatch", "SegmentListResult",
           "MediaMap", "SegmentView", "TrackView", "JSError", "JSTypeError", "undefined",
           "_native"]

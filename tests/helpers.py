"""shared helpers for the tests (test infrastructure: may import oracle/)"""
import hashlib
import json
import os

import numpy as np

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_host
from oracle import oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ALL_OPS = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX
SEED = 0xBE401DE5
SUM_RTOL = 1e-6   # north_star: floating-point sums within 1e-6 relative (reduction order)


def load_json(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def typed_schema():
    fields = [("mediaId", 1, L.K_STRING, L.R_KEY), ("status", 2, L.K_INT32, L.R_STATUS),
              ("progress", 3, L.K_DOUBLE, L.R_VALUE), ("host", 4, L.K_STRING, 0), ("blob", 5, L.K_BYTES, 0),
              ("delta", 6, L.K_SINT32, 0), ("ratio", 7, L.K_FLOAT, 0), ("flag", 8, L.K_BOOL, 0), ("n", 9, L.K_UINT32, 0)]
    return bh.Schema([bh.Field(*f) for f in fields], dict(bh.telemetry_progress().enum))


def pack(payloads):
    """list of bytes -> (slab u8 padded to 16, offsets u32[n+1])"""
    off = np.zeros(len(payloads) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(p) for p in payloads])
    raw = b"".join(payloads)
    slab = np.zeros((len(raw) + 15) // 16 * 16 + 16, dtype=np.uint8)
    slab[: len(raw)] = np.frombuffer(raw, dtype=np.uint8)
    return slab, off


def stream_digest(slab, off, ts):
    return hashlib.sha256(slab[: off[-1]].tobytes() + off.tobytes() + ts.tobytes()).hexdigest()


def sort_records(recs, hist=None):
    order = np.lexsort((recs["key_hash"], recs["window_id"]))
    return (recs[order], hist[order]) if hist is not None else recs[order]


def assert_records_equal(got, want, ops=ALL_OPS, got_hist=None, want_hist=None):
    """integer aggregates, keys, min/max and histogram buckets bit-exact; sums within SUM_RTOL"""
    if got_hist is not None:
        got, got_hist = sort_records(got, got_hist)
        want, want_hist = sort_records(want, want_hist)
    else:
        got, want = sort_records(got), sort_records(want)
    assert len(got) == len(want), (len(got), len(want))
    for f in ("window_id", "key_hash", "count"):
        assert np.array_equal(got[f], want[f]), f
    if ops & L.OP_MIN:
        assert np.array_equal(got["min"].view(np.uint64), want["min"].view(np.uint64)), "min"
    if ops & L.OP_MAX:
        assert np.array_equal(got["max"].view(np.uint64), want["max"].view(np.uint64)), "max"
    if ops & L.OP_SUM:
        np.testing.assert_allclose(got["sum"], want["sum"], rtol=SUM_RTOL, atol=0)
    if got_hist is not None:
        assert np.array_equal(got_hist, want_hist), "hist"

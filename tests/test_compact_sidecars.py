"""compact_sidecars (host side of bh_push_batch_compact): the u32 offsets and u64 timestamps of a batch can be
rebuilt exactly from the u16 lengths, i16 deltas and the per-block offsets / base timestamps -- the arithmetic the
GPU expansion kernel performs, restated in numpy.  No GPU."""
import numpy as np
import pytest

from beholder_b200 import _lib as L
from beholder_b200.engine import compact_sidecars
from beholder_b200.synth import SynthSpec, generate_host


def _rebuild(lens, boff, tsb, tsd):
    B = L.BH_COMPACT_BLOCK
    n = len(lens)
    off = np.zeros(n + 1, dtype=np.uint64)
    for k in range((n + B - 1) // B):
        lo, hi = k * B, min(n, (k + 1) * B)
        off[lo:hi] = boff[k] + np.concatenate([[0], np.cumsum(lens[lo:hi - 1].astype(np.uint64))]).astype(np.uint64)
    off[n] = boff[-1]
    ts = None
    if tsb is not None:
        ts = (np.repeat(tsb.astype(np.int64), B)[:n] + tsd.astype(np.int64)).astype(np.uint64)
    return off, ts


@pytest.mark.parametrize("n", [1, 999, 65_536, 65_537, 150_000])
def test_round_trip(n):
    slab, off, ts = generate_host(SynthSpec(n_keys=20, payload_bytes=0, rate_per_ms=3, jitter_ms=3000), 0, min(n, 3000))
    reps = -(-n // (len(off) - 1))
    lens0 = np.tile(np.diff(off.astype(np.int64)), reps)[:n]
    off2 = np.concatenate([[0], np.cumsum(lens0)]).astype(np.uint32)
    ts2 = (np.tile(ts.astype(np.int64) - int(ts[0]), reps)[:n] + 1_700_000_000_000 + np.arange(n) // 7).astype(np.uint64)
    lens, boff, tsb, tsd = compact_sidecars(off2, ts2)
    assert lens.dtype == np.uint16 and tsd.dtype == np.int16 and len(boff) == (n + L.BH_COMPACT_BLOCK - 1) // L.BH_COMPACT_BLOCK + 1
    r_off, r_ts = _rebuild(lens, boff, tsb, tsd)
    assert np.array_equal(r_off, off2.astype(np.uint64)) and np.array_equal(r_ts, ts2)
    # per event on the link: 2 + 2 bytes instead of 4 + 8
    assert lens.nbytes + tsd.nbytes + boff.nbytes + tsb.nbytes < 0.34 * (off2.nbytes + ts2.nbytes) + 64


def test_offsets_that_do_not_start_at_zero_and_no_timestamps():
    off = np.array([128, 192, 192, 300], dtype=np.uint32)
    lens, boff, tsb, tsd = compact_sidecars(off)
    assert lens.tolist() == [64, 0, 108] and boff.tolist() == [0, 172] and tsb is None and tsd is None

"""bh_push_batch_compact: u16 lengths + i16 timestamp deltas (+ one byte offset and one base timestamp per 65 536
messages) instead of u32 offsets + u64 timestamps -- 68 instead of 76 bytes per 64-byte event over the PCIe link,
which is what bounds the path end to end.  Offsets and timestamps are rebuilt on the GPU; results must be those of
the plain batch call, i.e. the oracle's."""
import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.engine import compact_sidecars
from beholder_b200.synth import SynthSpec, generate_host
from oracle import oracle as O
from helpers import ALL_OPS, SEED, assert_records_equal

pytestmark = pytest.mark.gpu
KEY = ("mediaId", "status")


def _pinned(*arrays):
    out = []
    for a in arrays:
        b = bh.PinnedBuffer(max(a.nbytes, 1) + 64)
        v = b.view(a.dtype, count=len(a))
        v[:] = a
        out.append((b, v))
    return out


@pytest.mark.parametrize("n,payload_bytes,slab_events", [(200_000, 0, 65_536), (131_072, 64, 65_536), (70_001, 64, 1 << 17), (999, 0, 1 << 16)])
def test_compact_batches_match_the_oracle(n, payload_bytes, slab_events):
    spec = SynthSpec(n_keys=500, rate_per_ms=20, jitter_ms=1500, payload_bytes=payload_bytes, malformed_ppm=3000, unknown_enum_ppm=3000)
    slab, off, ts = generate_host(spec, 0, n)
    lens, boff, tsb, tsd = compact_sidecars(off, ts)
    bufs = _pinned(slab, lens, boff, tsb, tsd)
    (_, hp), (_, hl), (_, hb), (_, htb), (_, htd) = bufs
    with bh.Aggregator(mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, table_capacity=1 << 12, slab_events=slab_events,
                       slab_bytes=8 << 20) as g:
        g.push_batch_compact(hp, hl, hb, htb, htd)
        g.push(bh.encode(bh.telemetry_progress(), mediaId="tail", status=1, progress=5), int(ts[-1]))     # ring path after it
        g.flush()
        got = g.snapshot_all()
        k = g.counters()
    a = O.Agg(bh.telemetry_progress().with_roles(key=KEY), ALL_OPS, 10_000, SEED)
    a.push_raw(slab, off, ts)
    tail = bh.encode(bh.telemetry_progress(), mediaId="tail", status=1, progress=5)
    a.push_raw(np.frombuffer(tail + b"\0" * 16, dtype=np.uint8), np.array([0, len(tail)], np.uint32), ts[-1:])
    assert_records_equal(got, a.snapshot(), ALL_OPS)
    d = a.dropped()
    assert (k["dropped_decode"], k["dropped_enum"]) == (d["decode"], d["enum"]) and k["events_ok"] == int(a.snapshot()["count"].sum())
    for b, _ in bufs:
        b.free()


def test_compact_reference_mode_without_timestamps():
    n = 100_000
    slab, off, ts = generate_host(SynthSpec(n_keys=50, payload_bytes=64, unknown_enum_ppm=5000), 0, n)
    lens, boff, _, _ = compact_sidecars(off)
    bufs = _pinned(slab, lens, boff)
    with bh.Aggregator(mode=L.MODE_REFERENCE) as g:
        g.push_batch_compact(bufs[0][1], bufs[1][1], bufs[2][1])
        g.flush()
        counts, k = g.status_counts(), g.counters()
    ref = O.Handler(bh.telemetry_progress())
    ref.handle_slab(slab, off)
    assert counts == ref.counts() and {"decode": k["dropped_decode"], "enum": k["dropped_enum"]} == ref.dropped()
    for b, _ in bufs:
        b.free()


def test_sidecars_that_do_not_fit_are_refused_on_the_host():
    off = np.array([0, 70_000], dtype=np.uint32)
    with pytest.raises(ValueError):
        compact_sidecars(off)
    off = np.arange(0, 64 * 11, 64, dtype=np.uint32)
    ts = np.array([0] * 9 + [100_000], dtype=np.uint64)
    with pytest.raises(ValueError):
        compact_sidecars(off, ts)
    with bh.Aggregator(mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, slab_events=1024) as g:
        with pytest.raises(bh.BeholderError) as e:
            g.push_batch_compact(np.zeros(64, np.uint8), np.array([64], np.uint16), np.array([0, 64], np.uint64),
                                 np.array([0], np.uint64), np.array([0], np.int16))
        assert e.value.code == L.BH_ERR_INVALID

"""A caller-owned (pinned) batch whose offsets table is damaged: the library does not scan the offsets on the host
(it would cost a tenth of the ingest rate), so the kernel must treat every offset outside the bytes that were
copied as a message that does not decode -- proto.decode would throw on such a buffer, index.js:129,149 -- and
read nothing outside them."""
import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_host
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def test_damaged_offsets_in_a_pinned_batch_are_dropped_not_followed():
    n = 50_000
    slab, off, ts = generate_host(SynthSpec(n_keys=50, payload_bytes=64), 0, n)
    bad = off.copy()
    rng = np.random.default_rng(3)
    hit = np.sort(rng.choice(np.arange(100, n - 100), 200, replace=False))
    for i in hit[:100]:
        bad[i] = np.uint32(0xFFFFFF00 - int(rng.integers(0, 1 << 20)))     # far beyond the slab
    for i in hit[100:]:
        bad[i] = np.uint32(int(rng.integers(0, 64)))                       # far below its neighbours
    pb = bh.PinnedBuffer(slab.nbytes + 64); ob = bh.PinnedBuffer(bad.nbytes); tb = bh.PinnedBuffer(ts.nbytes)
    hp = pb.view(np.uint8); hp[: len(slab)] = slab
    ho = ob.view(np.uint32); ho[:] = bad
    ht = tb.view(np.uint64); ht[:] = ts
    with bh.Aggregator(mode=L.MODE_REFERENCE) as g:
        g.push_batch(hp[: len(slab)], ho, None, pinned=True)
        g.flush()
        counts, k = g.status_counts(), g.counters()
    # a damaged offset spoils the message that ends there and the one that starts there; everything else counts
    spoiled = set(hit.tolist()) | set((hit - 1).tolist())
    keep = np.array([i for i in range(n) if i not in spoiled])
    ref = O.Handler(bh.telemetry_progress())
    for i in keep:
        ref.handle(bytes(slab[off[i]: off[i + 1]]))
    assert sum(counts.values()) >= sum(ref.counts().values())              # (a spoiled span may still parse as a message)
    assert k["events_ok"] + k["dropped_decode"] + k["dropped_enum"] == n
    assert k["dropped_decode"] >= len(spoiled) * 0.5
    for name, c in ref.counts().items():
        assert counts.get(name, 0) >= c
    pb.free(); ob.free(); tb.free()


def test_a_table_whose_end_is_below_its_start_copies_nothing_and_drops_everything():
    slab, off, ts = generate_host(SynthSpec(n_keys=5, payload_bytes=64), 0, 100)
    bad = off.copy(); bad[-1] = 0                       # the batch now claims to end where it starts
    pb = bh.PinnedBuffer(slab.nbytes + 64); ob = bh.PinnedBuffer(bad.nbytes)
    hp = pb.view(np.uint8); hp[: len(slab)] = slab
    ho = ob.view(np.uint32); ho[:] = bad
    with bh.Aggregator(mode=L.MODE_REFERENCE) as g:
        g.push_batch(hp[: len(slab)], ho, None, pinned=True)
        g.flush()
        k = g.counters()
    assert k["events_ok"] + k["dropped_decode"] + k["dropped_enum"] == 100 and k["dropped_decode"] >= 99
    pb.free(); ob.free()

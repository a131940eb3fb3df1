"""Full-key identity ("collisions are detected ..., never silently merged"; the reference keys
its counter by the label string itself, index.js:136-138, so it cannot collide).  With BH_F_VERIFY_KEYS the
library keeps the key bytes first seen with every 64-bit hash and compares every later event's key against them.
Real 64-bit collisions cannot be produced on demand, so the test hook `debug_hash_mask` truncates the hash."""
import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_device
from oracle import oracle as O
from helpers import ALL_OPS, SEED, assert_records_equal

pytestmark = pytest.mark.gpu
KEY = ("mediaId", "status")


def _stream(n=400_000, keys=3000, **kw):
    spec = SynthSpec(n_keys=keys, rate_per_ms=50, **kw)
    return generate_device(spec, 0, n)[:3]


@pytest.mark.parametrize("payload_bytes", [0, 64])
@pytest.mark.parametrize("flags", [0, L.F_PARTITIONED])
def test_no_collision_reported_on_distinct_hashes_and_results_unchanged(payload_bytes, flags):
    p, o, t = _stream(payload_bytes=payload_bytes, malformed_ppm=2000, unknown_enum_ppm=2000)
    n = len(t)
    with bh.Aggregator(mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, table_capacity=1 << 14, window_ring=8, verify_keys=True,
                       flags=flags) as g:
        for _ in range(2):                                  # the second pass finds every key in the dictionary
            g.process_device(p, o, t, n)
        g.flush()                                           # raises on BH_ERR_KEY_COLLISION
        got = g.snapshot_all()
        k = g.counters()
    assert k["key_collisions"] == 0
    slab = p.cpu().numpy(); off = o.cpu().numpy().view(np.uint32); ts = t.cpu().numpy().view(np.uint64)
    a = O.Agg(bh.telemetry_progress().with_roles(key=KEY), ALL_OPS, 10_000, SEED)
    a.push_raw(slab, off, ts); a.push_raw(slab, off, ts)
    assert a.hash_collisions() == 0
    assert_records_equal(got, a.snapshot(), ALL_OPS)


def test_truncated_hashes_collide_and_the_collision_is_reported():
    p, o, t = _stream(n=200_000, keys=3000, payload_bytes=64)
    n = len(t)
    with bh.Aggregator(mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, table_capacity=1 << 12, window_ring=8, verify_keys=True,
                       debug_hash_mask=0xFF) as g:         # 3000 keys onto 256 hash values
        g.process_device(p, o, t, n)
        with pytest.raises(bh.BeholderError) as e:
            g.flush()
        assert e.value.code == L.BH_ERR_KEY_COLLISION and "different key" in str(e.value)
        k = g.counters(check=False)
        rows = g.snapshot_all()
    # at most 256 keys can be "first seen": every event of the other ~2700 keys is reported
    assert 0.8 * n < k["key_collisions"] < n
    assert len(np.unique(rows["key_hash"])) <= 256
    # without the check the same stream is merged silently: that is what the flag is for
    with bh.Aggregator(mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, table_capacity=1 << 12, window_ring=8,
                       debug_hash_mask=0xFF) as g:
        g.process_device(p, o, t, n)
        g.flush()
        assert g.counters()["key_collisions"] == 0


def test_long_keys_and_the_ingest_path():
    # keys longer than the 40 bytes an entry stores are compared by their first 32 bytes and a hash of the rest;
    # messages go through bh_push_batch (the pinned ring), where the check runs on the ring slab
    schema = bh.telemetry_progress()
    msgs, tss = [], []
    rng = np.random.default_rng(5)
    keys = [("k%03d-" % i) + "x" * int(rng.integers(30, 90)) for i in range(300)]
    for i in range(60_000):
        msgs.append(bh.encode(schema, mediaId=keys[int(rng.integers(0, 300))], status=int(rng.integers(0, 6)), progress=int(rng.integers(0, 101))))
        tss.append(1_700_000_000_000 + i // 10)
    off = np.zeros(len(msgs) + 1, dtype=np.uint32); off[1:] = np.cumsum([len(m) for m in msgs])
    slab = np.frombuffer(b"".join(msgs) + b"\0" * 32, dtype=np.uint8).copy()
    ts = np.array(tss, dtype=np.uint64)
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, table_capacity=1 << 12, verify_keys=True,
                       message_bytes_hint=110) as g:
        g.push_batch(slab, off, ts)
        g.flush()
        assert g.counters()["key_collisions"] == 0
        got = g.snapshot_all()
    a = O.Agg(schema.with_roles(key=KEY), ALL_OPS, 10_000, SEED)
    a.push_raw(slab, off, ts)
    assert_records_equal(got, a.snapshot(), ALL_OPS)
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, table_capacity=1 << 12, verify_keys=True,
                       message_bytes_hint=110, debug_hash_mask=0x3F) as g:
        g.push_batch(slab, off, ts)
        with pytest.raises(bh.BeholderError) as e:
            g.flush()
        assert e.value.code == L.BH_ERR_KEY_COLLISION


def test_columns_and_unsupported_key_shapes_are_refused():
    with bh.Aggregator(mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, verify_keys=True) as g:
        kh = np.arange(10, dtype=np.uint64); ts = np.zeros(10, dtype=np.uint64); v = np.zeros(10)
        with pytest.raises(bh.BeholderError) as e:
            g.push_columns(kh, ts, v)
        assert e.value.code == L.BH_ERR_UNSUPPORTED
    with pytest.raises(bh.BeholderError) as e:
        bh.Aggregator(mode=L.MODE_WINDOWED, key=("status",), ops=ALL_OPS, verify_keys=True)
    assert e.value.code == L.BH_ERR_UNSUPPORTED

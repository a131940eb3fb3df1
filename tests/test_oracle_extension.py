"""Properties of the oracle's extension aggregator and handler (index.js:127-155)."""
import numpy as np

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_host
from oracle import oracle as O
from helpers import ALL_OPS, SEED, assert_records_equal


def test_threads_equal_single():
    spec = SynthSpec(n_keys=100, rate_per_ms=2, malformed_ppm=20000, unknown_enum_ppm=20000)
    slab, off, ts = generate_host(spec, 0, 6000)
    schema = bh.telemetry_progress().with_roles(key=("mediaId", "status"))
    one = O.Agg(schema, ALL_OPS | L.OP_HIST, 10_000, SEED)
    one.push_raw(slab, off, ts)
    many = O.Agg.run_threads(schema, ALL_OPS | L.OP_HIST, 10_000, SEED, slab, off, ts, 4)
    r1, h1 = one.snapshot(True)
    r2, h2 = many.snapshot(True)
    assert_records_equal(r2, r1, ALL_OPS, h2, h1)
    assert one.dropped() == many.dropped()
    h = O.Handler(bh.telemetry_progress())
    h.handle_slab(slab, off)
    counts, dropped = O.run_handler_threads(bh.telemetry_progress(), slab, off, 3)
    assert counts == h.counts()
    assert int(dropped[1]) == h.dropped()["decode"] and int(dropped[2]) == h.dropped()["enum"]


def test_logging_mode_does_not_change_counts():
    spec = SynthSpec(n_keys=10, malformed_ppm=50000)
    slab, off, ts = generate_host(spec, 0, 500)
    a, b = O.Handler(bh.telemetry_progress(), 0), O.Handler(bh.telemetry_progress(), 1)
    a.handle_slab(slab, off); b.handle_slab(slab, off)
    assert a.counts() == b.counts() and a.dropped() == b.dropped()


def test_metrics_text_format():
    h = O.Handler(bh.telemetry_progress())
    for st in (4, 4, 1):
        h.handle(bh.encode(bh.telemetry_progress(), mediaId="a", status=st, progress=1))
    txt = h.metrics_text().splitlines()
    assert txt[0].startswith("# HELP beholder_progress_updates_total ")
    assert txt[1] == "# TYPE beholder_progress_updates_total counter"
    assert txt[2:] == ['beholder_progress_updates_total{status="deployed"} 2',
                       'beholder_progress_updates_total{status="downloading"} 1']


def test_columns_match_raw():
    spec = SynthSpec(n_keys=64, rate_per_ms=1, value_f64=True, value_signed=True)
    slab, off, ts = generate_host(spec, 0, 3000)
    schema = bh.telemetry_progress(L.K_DOUBLE).with_roles(key=("mediaId", "status"))
    kh, val, st, oc = O.decode_columns(schema, ALL_OPS, SEED, slab, off)
    assert (oc == 0).all()
    raw = O.Agg(schema, ALL_OPS, 10_000, SEED); raw.push_raw(slab, off, ts)
    col = O.Agg(None, ALL_OPS, 10_000, SEED); col.push_columns(kh, ts, val)
    assert_records_equal(col.snapshot(), raw.snapshot())


def test_hist_bucket_and_order():
    B, E0 = 64, 991
    assert O.lib.bho_hist_bucket(0.0, B, E0) == 0
    assert O.lib.bho_hist_bucket(-0.0, B, E0) == 0
    assert O.lib.bho_hist_bucket(-3.0, B, E0) == 0
    assert O.lib.bho_hist_bucket(1.0, B, E0) == 32
    assert O.lib.bho_hist_bucket(1.999, B, E0) == 32
    assert O.lib.bho_hist_bucket(2.0, B, E0) == 33
    assert O.lib.bho_hist_bucket(5e-324, B, E0) == 1
    assert O.lib.bho_hist_bucket(float("inf"), B, E0) == 63
    assert O.lib.bho_hist_bucket(1e300, B, E0) == 63
    vals = [-np.inf, -1e9, -1.0, -5e-324, -0.0, 0.0, 5e-324, 1.0, 1e9, np.inf]
    ords = [O.lib.bho_f64_order(v) for v in vals]
    assert ords == sorted(ords) and len(set(ords)) == len(ords)


def test_key_hash_has_no_collisions_on_many_uuid_keys():
    # 200k distinct synthetic (mediaId, status) keys, one event each
    spec = SynthSpec(n_keys=1 << 62, rate_per_ms=1000, jitter_ms=0, payload_bytes=0)
    slab, off, ts = generate_host(spec, 0, 20000)
    schema = bh.telemetry_progress().with_roles(key=("mediaId", "status"))
    a = O.Agg(schema, L.OP_COUNT, 0, SEED)
    a.push_raw(slab, off, ts)
    assert a.hash_collisions() == 0
    recs = a.snapshot()
    assert len(np.unique(recs["key_hash"])) == len(recs)
    # low bits and high bits both spread (table index uses high bits of kh * phi; owner uses top bits)
    assert len(np.unique(recs["key_hash"] >> np.uint64(56))) == 256
    assert len(np.unique(recs["key_hash"] & np.uint64(0xFF))) == 256


def test_key_hash_separates_keys_that_differ_in_a_few_chosen_bits():
    """A multiply leaves a difference in bit 63 of its operand as a difference in bit 63 only, which the next
    8-byte word can cancel.  Earlier versions of the hash collided on keys differing in {bit 63 of word i,
    bit 31 of word i+1} and then in three such bits.  All 36-byte keys within one, two or (for the bits a
    multiply treats specially: 15, 31, 47, 63 of every word) three flipped bits of each other, plus 20 000
    random flip sets of up to 6 bits, must hash apart."""
    import ctypes as C
    import itertools
    rng = np.random.default_rng(SEED)
    base = bytearray(b"d418035b-9722-015d-8d9a-e76ed8cd1872")
    nbits = len(base) * 8

    def h(b):
        buf = (C.c_uint8 * len(b)).from_buffer_copy(bytes(b))
        return O.lib.bho_hash_bytes(C.addressof(buf), len(b), 1, SEED)

    def flipped(bits):
        b = bytearray(base)
        for i in bits:
            b[i // 8] ^= 1 << (i % 8)
        return b

    sets = {()}
    sets |= {(i,) for i in range(nbits)}
    sets |= set(itertools.combinations(range(nbits), 2))
    special = [w * 64 + b for w in range(nbits // 64 + 1) for b in (15, 31, 47, 63) if w * 64 + b < nbits]
    sets |= set(itertools.combinations(special, 3))
    for _ in range(20000):
        k = int(rng.integers(3, 7))
        sets.add(tuple(sorted(rng.choice(nbits, k, replace=False).tolist())))
    seen = {}
    for p in sets:
        v = h(flipped(p))
        assert v not in seen, (p, seen[v])
        seen[v] = p

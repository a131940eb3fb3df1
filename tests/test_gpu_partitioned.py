"""GPU parity of the partitioned (deferred) aggregation used for large key sets — BASELINE.json configs 3 and 4
(1 M / 10 M keys; no counterpart in the reference, whose counter has a handful of labels, index.js:136-138).
Events that miss the per-SM shared-memory table are logged as tuples partitioned by table slot and folded into
the HBM tables later; the result must be the oracle's on the same seeded stream: keys, counts, min/max and
histogram buckets bit-exact, f64 sums within 1e-6 relative."""
import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_device, columns_device
from oracle import oracle as O
from helpers import ALL_OPS, SEED, assert_records_equal

pytestmark = pytest.mark.gpu
KEY = ("mediaId", "status")


def _stream(spec, first, n):
    p, o, t, _ = generate_device(spec, first, n)
    return p, o, t


def _check(spec, n_batches, per_batch, ops, cap, *, flags=L.F_PARTITIONED, ring=4, threads=32, reads_between=False):
    schema = bh.telemetry_progress(L.K_DOUBLE if spec.value_f64 else L.K_INT32)
    a = O.Agg(schema.with_roles(key=KEY), ops, 10_000, SEED)
    hist = bool(ops & L.OP_HIST)
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ops, window_ms=10_000, table_capacity=cap,
                       window_ring=ring, flags=flags) as g:
        for b in range(n_batches):
            p, o, t = _stream(spec, b * per_batch, per_batch)
            g.process_device(p, o, t, per_batch)
            if reads_between and b == 0:
                g.counters()                       # a read folds the log; later batches start a new epoch
            slab = p.cpu().numpy(); off = o.cpu().numpy().view(np.uint32); ts = t.cpu().numpy().view(np.uint64)
            part = O.Agg.run_threads(schema.with_roles(key=KEY), ops, 10_000, SEED, slab, off, ts, threads)
            a.merge(part)
        g.sync()
        got = g.snapshot_all(with_hist=hist)
        k = g.counters()
    assert a.hash_collisions() == 0
    if hist:
        want, want_hist = a.snapshot(with_hist=True)
        assert_records_equal(got[0], want, ops, got[1], want_hist)
        assert int(got[1].sum()) == int(want["count"].sum())
    else:
        want = a.snapshot()
        assert_records_equal(got, want, ops)
    d = a.dropped()
    assert (k["dropped_decode"], k["dropped_enum"], k["dropped_nan"]) == (d["decode"], d["enum"], d["nan"])
    assert k["events_ok"] == int(want["count"].sum())
    assert k["dropped_table_full"] == 0 and k["dropped_window"] == 0
    return k


@pytest.mark.parametrize("zipf", [False, True])
@pytest.mark.parametrize("f64", [False, True])
def test_forced_on_a_small_table(zipf, f64):
    # 20 k keys, ~10 k events per SM and batch against 4 096 shared slots: many events are logged; 3 batches, a
    # read after the first
    spec = SynthSpec(n_keys=20_000, rate_per_ms=200, zipf=zipf, value_f64=f64, value_signed=f64,
                     malformed_ppm=3000, unknown_enum_ppm=3000)
    ops = ALL_OPS | (L.OP_HIST if f64 else 0)
    k = _check(spec, 3, 1_500_000, ops, 1 << 16, reads_between=True, threads=16)
    assert k["smem_spills"] > (100_000 if zipf else 1_000_000)


@pytest.mark.parametrize("f64", [False, True])
@pytest.mark.parametrize("ops", [L.OP_COUNT, L.OP_COUNT | L.OP_MIN, L.OP_COUNT | L.OP_MAX, L.OP_SUM, L.OP_SUM | L.OP_MIN | L.OP_MAX])
def test_single_event_entries_logged_at_the_flush_under_every_op_subset(ops, f64):
    # many keys, few events each, no histograms: most front-table entries end a launch holding ONE event and are
    # logged as a tuple whose value comes from whichever column the ops keep
    spec = SynthSpec(n_keys=400_000, rate_per_ms=200, value_f64=f64, value_signed=True)
    _check(spec, 2, 600_000, ops, 1 << 20, threads=16)


@pytest.mark.parametrize("sub_lg", ["30", "7"])
@pytest.mark.parametrize("f64", [False, True])
def test_slice_mirror_fold_forced(monkeypatch, f64, sub_lg):
    # the fold that holds a partition's slice of the table in shared memory (by itself it is used from 8192-slot
    # slices on): whole slices, and 128-slot parts of 256-slot slices so that probe chains run off the part held in
    # shared memory and go on in the HBM table; with histograms (16-bit counters, several rounds per partition)
    monkeypatch.setenv("BEHOLDER_FOLD", "1")
    monkeypatch.setenv("BEHOLDER_FOLD_SUB_LG", sub_lg)
    spec = SynthSpec(n_keys=20_000, rate_per_ms=200, value_f64=f64, value_signed=f64, malformed_ppm=3000, unknown_enum_ppm=3000)
    k = _check(spec, 3, 1_500_000, ALL_OPS | (L.OP_HIST if f64 else 0), 1 << 16, reads_between=True, threads=16)
    assert k["smem_spills"] > 1_000_000
    # a well filled table (load 0.73, 256-slot slices holding 188 +- 14 keys): long chains, wrapping inside the slices
    spec = SynthSpec(n_keys=6_000, rate_per_ms=200, zipf=True, value_f64=f64)
    _check(spec, 2, 1_000_000, ALL_OPS, 1 << 13, threads=8)


def test_slice_mirror_fold_at_the_size_that_selects_it():
    # 2^24 slots -> 2048 partitions of 8192 slots, folded in two 4096-slot parts; 3 M keys, ~2 events per key
    spec = SynthSpec(n_keys=3_000_000, rate_per_ms=1000)
    k = _check(spec, 2, 3_000_000, ALL_OPS, 1 << 24, flags=0, ring=4)
    assert k["smem_spills"] > 4_000_000


def test_forcing_it_on_a_table_too_small_for_256_slot_slices_is_refused():
    for cap, ring in ((256, 4), (512, 4), (4096, 32)):
        with pytest.raises(bh.BeholderError) as e:
            bh.Aggregator(bh.telemetry_progress(), mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, table_capacity=cap,
                          window_ring=ring, flags=L.F_PARTITIONED)
        assert e.value.code == L.BH_ERR_INVALID
    bh.Aggregator(bh.telemetry_progress(), mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, table_capacity=1024, window_ring=4,
                  flags=L.F_PARTITIONED).close()


def test_partitioning_is_on_by_itself_for_large_tables_and_can_be_kept_off():
    spec = SynthSpec(n_keys=100_000, rate_per_ms=400)
    k_on = _check(spec, 1, 2_000_000, ALL_OPS, 1 << 18, flags=0, threads=16)
    k_off = _check(spec, 1, 2_000_000, ALL_OPS, 1 << 18, flags=L.F_NO_PARTITION, threads=16)
    assert k_on["events_ok"] == k_off["events_ok"] and k_on["smem_spills"] > 500_000


def test_skewed_keys_overflow_their_log_regions_and_still_match():
    # Zipf over few keys with a tiny table in front: the hottest partitions overflow their regions, the
    # overflow takes the direct path; nothing may be lost or counted twice
    spec = SynthSpec(n_keys=3000, rate_per_ms=200, zipf=True)
    _check(spec, 2, 1_500_000, ALL_OPS, 1 << 13, threads=8)


def test_columns_and_raw_events_in_one_log():
    # an int32 schema fed through both entry points: the drain must treat the logged values as f64
    import torch
    spec = SynthSpec(n_keys=30_000, rate_per_ms=50)
    n = 400_000
    schema = bh.telemetry_progress()
    a = O.Agg(schema.with_roles(key=KEY), ALL_OPS, 10_000, SEED)       # groups by key bytes
    b = O.Agg(None, ALL_OPS, 10_000, SEED)                              # groups by the given key hash
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, table_capacity=1 << 16, window_ring=4,
                       flags=L.F_PARTITIONED) as g:
        p, o, t = _stream(spec, 0, n)
        g.process_device(p, o, t, n)
        a.push_raw(p.cpu().numpy(), o.cpu().numpy().view(np.uint32), t.cpu().numpy().view(np.uint64))
        kh, ts, val = columns_device(spec, SEED, n, n)
        val = val + 0.25                                   # not integers
        g.process_columns_device(kh, ts, val, n)
        b.push_columns(kh.cpu().numpy().view(np.uint64), ts.cpu().numpy().view(np.uint64), val.cpu().numpy())
        g.sync()
        got = g.snapshot_all()
    # the two oracle tables joined on (window, key hash)
    rows = {}
    for r in list(a.snapshot()) + list(b.snapshot()):
        k = (int(r["window_id"]), int(r["key_hash"]))
        if k in rows:
            q = rows[k]
            rows[k] = (q[0] + int(r["count"]), q[1] + float(r["sum"]), min(q[2], float(r["min"])), max(q[3], float(r["max"])))
        else:
            rows[k] = (int(r["count"]), float(r["sum"]), float(r["min"]), float(r["max"]))
    want = np.zeros(len(rows), dtype=got.dtype)
    for i, ((w, h), (c, s_, mn, mx)) in enumerate(rows.items()):
        want[i] = (h, w, c, s_, mn, mx)
    assert_records_equal(got, want, ALL_OPS)


@pytest.mark.parametrize("zipf", [False, True])
def test_config3_shape_one_million_keys_with_histograms(zipf):
    # BASELINE.json configs[2] at reduced length: 1 M keys, 64-bucket log2 histograms, f64 values, 24 M events in
    # 4 batches (the oracle's threaded port aggregates each batch, partial tables are merged)
    spec = SynthSpec(n_keys=1_000_000, rate_per_ms=2000, zipf=zipf, value_f64=True, value_exp_lo=-12, value_exp_span=24)
    k = _check(spec, 4, 6_000_000, ALL_OPS | L.OP_HIST, 1 << 21, flags=0, ring=4)
    assert k["smem_spills"] > 6_000_000

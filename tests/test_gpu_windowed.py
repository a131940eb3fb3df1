"""GPU parity, windowed extension: keys/counts/min/max/histograms bit-exact,
f64 sums within 1e-6 relative, against the oracle on the same seeded streams."""
import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_host, generate_device, columns_device, zipf_cdf
from oracle import oracle as O
from helpers import ALL_OPS, SEED, assert_records_equal, load_json, pack, sort_records

pytestmark = pytest.mark.gpu
KEY = ("mediaId", "status")


def _oracle(schema, ops, window_ms, slab, off, ts, threads=8):
    a = O.Agg.run_threads(schema.with_roles(key=KEY), ops, window_ms, SEED, slab, off, ts, threads)
    return a


def _run_raw(spec, n, ops, window_ms=10_000, table_capacity=4096, first=0, **kw):
    import torch
    p, o, t, total = generate_device(spec, first, n)
    schema = bh.telemetry_progress(L.K_DOUBLE if spec.value_f64 else L.K_INT32)
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ops, window_ms=window_ms,
                       table_capacity=table_capacity, **kw) as g:
        g.process_device(p, o, t, n)
        g.sync()
        got = g.snapshot_all(with_hist=bool(ops & L.OP_HIST))
        k = g.counters(check=False)
    slab = p.cpu().numpy(); off = o.cpu().numpy().view(np.uint32); ts = t.cpu().numpy().view(np.uint64)
    a = _oracle(schema, ops, window_ms, slab, off, ts)
    assert a.hash_collisions() == 0
    return got, a, k


@pytest.mark.parametrize("zipf", [False, True])
@pytest.mark.parametrize("f64", [False, True])
def test_c2_shape_count_sum_min_max(zipf, f64):
    # BASELINE.json config 2 at reduced length: 1k keys, 10 s windows, count/sum/min/max
    spec = SynthSpec(n_keys=1000, rate_per_ms=10, zipf=zipf, value_f64=f64, value_signed=f64,
                     malformed_ppm=5000, unknown_enum_ppm=5000)
    got, a, k = _run_raw(spec, 1_000_000, ALL_OPS)
    assert_records_equal(got, a.snapshot(), ALL_OPS)
    d = a.dropped()
    assert (k["dropped_decode"], k["dropped_enum"], k["dropped_nan"]) == (d["decode"], d["enum"], d["nan"])
    assert k["events_ok"] == int(a.snapshot()["count"].sum())
    assert len(np.unique(got["window_id"])) >= 10


@pytest.mark.parametrize("generic", [False, True])
@pytest.mark.parametrize("zipf", [False, True])
def test_schema_specialised_and_generic_kernels_agree(generic, zipf):
    # int32 progress keyed by (mediaId, status) is the schema one fused kernel has compiled in; the
    # BH_F_GENERIC_KERNELS flag keeps it out.  Both must match the oracle on a stream with malformed
    # messages, unknown enum numbers, padded (host) and unpadded messages.
    for payload in (0, 64):
        spec = SynthSpec(n_keys=700, rate_per_ms=10, zipf=zipf, payload_bytes=payload,
                         malformed_ppm=20_000, unknown_enum_ppm=20_000)
        got, a, k = _run_raw(spec, 300_000, ALL_OPS, generic_kernels=generic)
        assert_records_equal(got, a.snapshot(), ALL_OPS)
        d = a.dropped()
        assert (k["dropped_decode"], k["dropped_enum"], k["dropped_nan"]) == (d["decode"], d["enum"], d["nan"])
        assert k["events_ok"] == int(a.snapshot()["count"].sum())


@pytest.mark.parametrize("generic", [False, True])
def test_kat_messages_windowed_both_kernels(generic):
    # the decoder known-answer messages (odd wire types, groups, unknown fields, truncations) as one
    # batch, repeated so that several warps see them, through the windowed path
    payloads = [bytes.fromhex(c["payload"]) for c in load_json("kat_decode.json")["default"]] * 40
    slab, off = pack(payloads)
    ts = (np.arange(len(payloads), dtype=np.uint64) * 3) + 1_700_000_000_000
    schema = bh.telemetry_progress()
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, window_ms=1000, table_capacity=256,
                       generic_kernels=generic) as g:
        g.push_batch(slab, off, ts)
        g.flush()
        got = g.snapshot_all()
        k = g.counters(check=False)
    a = _oracle(schema, ALL_OPS, 1000, slab, off, ts, threads=1)
    assert_records_equal(got, a.snapshot(), ALL_OPS)
    d = a.dropped()
    assert (k["dropped_decode"], k["dropped_enum"]) == (d["decode"], d["enum"])


@pytest.mark.parametrize("payload_bytes,hint", [(100, 100), (160, 160), (300, 300), (160, 0)])
def test_larger_messages_windowed(payload_bytes, hint):
    spec = SynthSpec(n_keys=500, rate_per_ms=10, payload_bytes=payload_bytes, value_f64=True, malformed_ppm=5000)
    got, a, k = _run_raw(spec, 150_000, ALL_OPS, table_capacity=1024, message_bytes_hint=hint)
    assert_records_equal(got, a.snapshot(), ALL_OPS)
    d = a.dropped()
    assert (k["dropped_decode"], k["dropped_enum"]) == (d["decode"], d["enum"])


def test_histograms_and_spill_many_keys():
    # config 3 shape at reduced size: more keys than the shared-memory table holds, log2 histograms
    ops = ALL_OPS | L.OP_HIST
    spec = SynthSpec(n_keys=50_000, rate_per_ms=100, value_f64=True, value_exp_lo=-40, value_exp_span=80)
    got, a, k = _run_raw(spec, 400_000, ops, table_capacity=1 << 17, window_ring=4)
    want, want_hist = a.snapshot(with_hist=True)
    assert_records_equal(got[0], want, ops, got[1], want_hist)
    assert k["smem_spills"] > 0
    assert got[1].sum() == want["count"].sum()


@pytest.mark.parametrize("window_ms", [1, 7, 1000, 10_000, 4096, 86_400_000])
def test_window_sizes(window_ms):
    # 667 ms of stream +- 500 ms jitter: ~240 windows of 7 ms (fits the ring), ~1700 of 1 ms (does not)
    spec = SynthSpec(n_keys=50, rate_per_ms=30, jitter_ms=500)
    got, a, k = _run_raw(spec, 20_000, L.OP_COUNT | L.OP_SUM, window_ms=window_ms, table_capacity=256,
                         window_ring=1024 if window_ms < 100 else 32)
    if k["dropped_window"] == 0:
        assert_records_equal(got, a.snapshot(), L.OP_COUNT | L.OP_SUM)
    else:   # window_ms=1: more windows than the ring holds must be REPORTED, never silent
        assert window_ms == 1


def test_columns_path_matches_raw_path_and_oracle():
    import torch
    spec = SynthSpec(n_keys=777, rate_per_ms=20, value_f64=True, value_signed=True, zipf=True)
    n = 500_000
    kh, ts, val = columns_device(spec, SEED, 0, n)
    with bh.Aggregator(mode=L.MODE_WINDOWED, ops=ALL_OPS, table_capacity=2048) as g:
        g.process_columns_device(kh, ts, val, n)
        g.sync()
        got_cols = g.snapshot_all()
    a = O.Agg(None, ALL_OPS, 10_000, SEED)
    a.push_columns(kh.cpu().numpy().view(np.uint64), ts.cpu().numpy().view(np.uint64), val.cpu().numpy())
    assert_records_equal(got_cols, a.snapshot())
    got_raw, a_raw, _ = _run_raw(spec, n, ALL_OPS, table_capacity=2048)
    assert_records_equal(got_raw, got_cols)


def test_push_columns_host_path_with_special_keys_and_nan():
    rng = np.random.default_rng(3)
    n = 100_000
    kh = rng.integers(0, 300, n).astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)
    kh[::1000] = np.uint64(0xFFFFFFFFFFFFFFFF)      # the table's EMPTY sentinel as a key
    kh[1::1000] = np.uint64(0xFFFFFFFFFFFFFFFE)     # the shared table's LOCK sentinel as a key
    kh[2::1000] = 0
    ts = (np.uint64(1_700_000_000_000) + rng.integers(0, 60_000, n).astype(np.uint64))
    val = rng.normal(0, 1e3, n)
    val[5::777] = np.nan
    val[6::777] = -0.0
    val[7::777] = np.inf
    ops = ALL_OPS | L.OP_HIST
    with bh.Aggregator(mode=L.MODE_WINDOWED, ops=ops, table_capacity=1024, slab_bytes=24 * 7000) as g:
        g.push_columns(kh, ts, val)
        g.flush()
        got, got_hist = g.snapshot_all(with_hist=True)
        k = g.counters()
    a = O.Agg(None, ops, 10_000, SEED)
    a.push_columns(kh, ts, val)
    want, want_hist = a.snapshot(with_hist=True)
    # inf sums: compare finite rows with tolerance, infinite rows exactly
    fin = np.isfinite(want["sum"])
    got_s, gh = (lambda r, h: (r[np.lexsort((r["key_hash"], r["window_id"]))], h[np.lexsort((r["key_hash"], r["window_id"]))]))(got, got_hist)
    assert np.array_equal(got_s["sum"][~fin], want["sum"][~fin])
    got_s["sum"][~fin] = 0; want = want.copy(); want["sum"][~fin] = 0
    assert_records_equal(got_s, want, ops, gh, want_hist)
    assert k["dropped_nan"] == a.dropped()["nan"] > 0


def test_table_full_is_an_error():
    spec = SynthSpec(n_keys=5000, rate_per_ms=1000, jitter_ms=0)
    import torch
    p, o, t, _ = generate_device(spec, 0, 50_000)
    with bh.Aggregator(mode=L.MODE_WINDOWED, ops=L.OP_COUNT, table_capacity=64) as g:
        g.process_device(p, o, t, 50_000)
        with pytest.raises(bh.BeholderError) as e:
            g.flush()
        assert e.value.code == L.BH_ERR_TABLE_FULL
        k = g.counters(check=False)
        assert k["dropped_table_full"] > 0 and k["dropped_table_full"] + k["events_ok"] == 50_000


def test_close_window_recycles_ring_slot():
    spec = SynthSpec(n_keys=20, rate_per_ms=1, jitter_ms=0)
    slab, off, ts = generate_host(spec, 0, 40_000)    # 40 s of stream -> windows w0..w0+4
    with bh.Aggregator(mode=L.MODE_WINDOWED, ops=L.OP_COUNT, table_capacity=64, window_ring=2,
                       slab_events=5000) as g:
        seen = {}
        for c in range(0, 40_000, 5000):
            sl = slice(c, c + 5000)
            o = off[c: c + 5001]
            g.push_batch(slab, o, ts[sl])
            g.flush()
            wins = g.list_windows()
            assert len(wins) <= 2
            for w in wins[:-1]:              # everything but the newest window is complete (no jitter)
                seen[w] = g.snapshot(w)
                g.close_window(w)
        for w in g.list_windows():
            seen[w] = g.snapshot(w)
    a = O.Agg(bh.telemetry_progress().with_roles(key=KEY), L.OP_COUNT, 10_000, SEED)
    a.push_raw(slab, off, ts)
    got = np.concatenate(list(seen.values()))
    assert_records_equal(got, a.snapshot(), L.OP_COUNT)


def test_decode_kernel_columns_match_oracle():
    import torch
    spec = SynthSpec(n_keys=300, value_f64=True, value_signed=True, malformed_ppm=50000, unknown_enum_ppm=50000,
                     payload_bytes=0)
    n = 100_000
    p, o, t, _ = generate_device(spec, 0, n)
    schema = bh.telemetry_progress(L.K_DOUBLE)
    dev = p.device
    kh = torch.empty(n, dtype=torch.int64, device=dev); val = torch.empty(n, dtype=torch.float64, device=dev)
    st = torch.empty(n, dtype=torch.int32, device=dev); oc = torch.empty(n, dtype=torch.uint8, device=dev)
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS) as g:
        g.decode_device(p, o, n, kh, val, st, oc)
        g.sync()
    wkh, wval, wst, woc = O.decode_columns(schema.with_roles(key=KEY), ALL_OPS, SEED,
                                           p.cpu().numpy(), o.cpu().numpy().view(np.uint32))
    assert np.array_equal(oc.cpu().numpy(), woc)
    assert np.array_equal(kh.cpu().numpy().view(np.uint64), wkh)
    assert np.array_equal(val.cpu().numpy().view(np.uint64), wval.view(np.uint64))
    assert np.array_equal(st.cpu().numpy(), wst)
    assert (woc != 0).sum() > 1000


def test_full_size_properties_c2():
    """BASELINE.json config 2 at full size (100 M events in 8 slabs), checked through
    size-independent properties: total count, linearity in the number of passes, idempotent
    min/max, and equality of raw vs columnar keys/counts."""
    import torch
    spec = SynthSpec(n_keys=1000, rate_per_ms=500)
    n_slab, slabs = 12_500_000, 8
    with bh.Aggregator(mode=L.MODE_WINDOWED, ops=ALL_OPS, table_capacity=2048, window_ring=32) as g, \
         bh.Aggregator(mode=L.MODE_WINDOWED, ops=ALL_OPS, table_capacity=2048, window_ring=32) as g2:
        buf = None
        for s in range(slabs):
            p, o, t, total = generate_device(spec, s * n_slab, n_slab, out=buf)
            buf = (p, o, t)
            g.process_device(p, o, t, n_slab)
            g2.process_device(p, o, t, n_slab); g2.process_device(p, o, t, n_slab)   # every event twice
            g.sync(); g2.sync()
        once, twice = g.snapshot_all(), g2.snapshot_all()
        k = g.counters()
    assert k["events_ok"] == n_slab * slabs == int(once["count"].sum())
    o1 = once[np.lexsort((once["key_hash"], once["window_id"]))]
    o2 = twice[np.lexsort((twice["key_hash"], twice["window_id"]))]
    assert np.array_equal(o1["key_hash"], o2["key_hash"]) and np.array_equal(o1["window_id"], o2["window_id"])
    assert np.array_equal(o2["count"], 2 * o1["count"])                   # linearity
    assert np.array_equal(o1["min"], o2["min"]) and np.array_equal(o1["max"], o2["max"])   # idempotence
    np.testing.assert_allclose(o2["sum"], 2 * o1["sum"], rtol=1e-6)
    assert len(np.unique(o1["window_id"])) >= 20 and len(np.unique(o1["key_hash"])) == 1000
    # integer progress 0..100: sums are integers and exactly representable -> must be exact
    assert np.array_equal(o1["sum"], np.round(o1["sum"])) and (o1["min"] >= 0).all() and (o1["max"] <= 100).all()


def test_integer_table_carries_and_negative_values():
    """int32 `progress` over the whole range: the shared table keeps the sum in two 32-bit words
    (native atomics, carry / sign handled by hand) and min/max as int32; multi-byte and 10-byte varints
    take the decoder's slow tiers.  Everything must equal the oracle exactly (sums are integers < 2^53)."""
    rng = np.random.default_rng(5)
    schema = bh.telemetry_progress()
    n = 60_000
    vals = np.concatenate([rng.integers(-2**31, 2**31, n // 2), rng.integers(-3, 130, n // 4),
                           np.full(n // 8, 2**31 - 1), np.full(n // 8, -2**31)]).astype(np.int64)
    rng.shuffle(vals)
    payloads = [bh.encode(schema, mediaId="m%d" % (i % 7), status=int(i % 6), progress=int(v)) for i, v in enumerate(vals)]
    slab, off = pack(payloads)
    ts = (np.uint64(1_700_000_000_000) + (np.arange(len(payloads), dtype=np.uint64) // np.uint64(3)))
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, table_capacity=256) as g:
        g.push_batch(slab, off, ts)
        g.flush()
        got = g.snapshot_all()
        k = g.counters()
    a = O.Agg(schema.with_roles(key=KEY), ALL_OPS, 10_000, SEED)
    a.push_raw(slab, off, ts)
    want = a.snapshot()
    assert_records_equal(got, want, ALL_OPS)
    gs, ws = sort_records(got), sort_records(want)
    assert np.array_equal(gs["sum"], ws["sum"])            # exact, not just within tolerance
    assert k["slow_decodes"] > n // 2 and k["events_ok"] == len(payloads)

"""GPU parity on adversarial input: valid progress messages mutated at random (bit flips, truncation,
inserted / appended fields of every wire type, groups, long varints, reordered and repeated fields,
pure noise), message sizes across all staging modes.  Every decoder tier and both fused kernels must
agree with the oracle event by event and in the aggregate."""
import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_host
from oracle import oracle as O
from helpers import ALL_OPS, SEED, assert_records_equal, pack

pytestmark = pytest.mark.gpu
KEY = ("mediaId", "status")


def _varint(v):
    out = bytearray()
    while True:
        b = v & 0x7F; v >>= 7
        out.append(b | (0x80 if v else 0))
        if not v:
            return bytes(out)


def _random_field(rng):
    num = int(rng.choice([1, 2, 3, 4, 5, 9, 15, 16, 31, 100]))
    wt = int(rng.integers(0, 8))
    tag = _varint(num << 3 | wt)
    if wt == 0:
        return tag + _varint(int(rng.integers(0, 1 << 63)) if rng.random() < 0.3 else int(rng.integers(0, 300)))
    if wt == 1:
        return tag + rng.bytes(8)
    if wt == 2:
        n = int(rng.integers(0, 40))
        return tag + _varint(n) + rng.bytes(n)
    if wt == 3:   # group: nested field(s) then the end-group tag
        inner = _random_field(rng) if rng.random() < 0.5 else b""
        return tag + inner + (_varint(num << 3 | 4) if rng.random() < 0.8 else b"")
    if wt == 5:
        return tag + rng.bytes(4)
    return tag      # 4 (stray end group), 6, 7 (invalid wire types)


def _mutate(msg, rng):
    m = bytearray(msg)
    r = rng.random()
    if r < 0.30:
        return bytes(m)
    if r < 0.40 and m:
        for _ in range(int(rng.integers(1, 4))):
            i = int(rng.integers(0, len(m))); m[i] ^= 1 << int(rng.integers(0, 8))
        return bytes(m)
    if r < 0.50:
        return bytes(m[: int(rng.integers(0, len(m) + 1))])
    if r < 0.60 and m:
        i = int(rng.integers(0, len(m))); m[i] = int(rng.integers(0, 256))
        return bytes(m)
    if r < 0.72:
        return bytes(m) + _random_field(rng)
    if r < 0.80:
        return _random_field(rng) + bytes(m)
    if r < 0.86:
        i = int(rng.integers(0, len(m) + 1))
        return bytes(m[:i]) + rng.bytes(int(rng.integers(1, 9))) + bytes(m[i:])
    if r < 0.92:
        return bytes(m) + bytes(m[int(rng.integers(0, len(m) + 1)):])          # repeated tail (duplicate fields)
    if r < 0.96:
        return rng.bytes(int(rng.integers(0, 200)))
    return bytes(m) + _varint(4 << 3 | 2) + _varint(400) + rng.bytes(int(rng.integers(0, 400)))   # long / overrunning string


def _fuzz_stream(f64, payload_bytes, n, seed):
    spec = SynthSpec(n_keys=200, value_f64=f64, value_signed=f64, payload_bytes=payload_bytes, rate_per_ms=20)
    slab, off, ts = generate_host(spec, 0, n)
    rng = np.random.default_rng(seed)
    msgs = [_mutate(bytes(slab[off[i]:off[i + 1]]), rng) for i in range(n)]
    s2, o2 = pack(msgs)
    return s2, o2, ts


@pytest.mark.parametrize("f64", [False, True])
@pytest.mark.parametrize("payload_bytes,hint", [(0, 0), (64, 0), (120, 120), (250, 250), (250, 0), (700, 700)])
def test_decoder_tiers_agree_with_oracle_on_mutated_messages(f64, payload_bytes, hint):
    import torch
    n = 30_000
    slab, off, ts = _fuzz_stream(f64, payload_bytes, n, SEED + payload_bytes + hint + int(f64))
    schema = bh.telemetry_progress(L.K_DOUBLE if f64 else L.K_INT32)
    p = torch.from_numpy(np.concatenate([slab, np.zeros(64, np.uint8)])).cuda()
    o = torch.from_numpy(off.view(np.int32)).cuda()
    kh = torch.empty(n, dtype=torch.int64, device="cuda"); val = torch.empty(n, dtype=torch.float64, device="cuda")
    st = torch.empty(n, dtype=torch.int32, device="cuda"); oc = torch.empty(n, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, message_bytes_hint=hint) as g:
        g.decode_device(p, o, n, kh, val, st, oc)
        g.sync()
    wkh, wval, wst, woc = O.decode_columns(schema.with_roles(key=KEY), ALL_OPS, SEED, slab, off)
    goc = oc.cpu().numpy()
    bad = np.nonzero(goc != woc)[0]
    assert bad.size == 0, (bad[:5], [bytes(slab[off[i]:off[i + 1]]).hex() for i in bad[:3]])
    assert np.array_equal(kh.cpu().numpy().view(np.uint64), wkh)
    assert np.array_equal(val.cpu().numpy().view(np.uint64), wval.view(np.uint64))
    assert np.array_equal(st.cpu().numpy(), wst)
    assert (woc == 0).sum() > n // 4 and (woc != 0).sum() > n // 20      # both kinds well represented


@pytest.mark.parametrize("generic", [False, True])
@pytest.mark.parametrize("f64", [False, True])
def test_fused_kernels_aggregate_mutated_stream_like_the_oracle(generic, f64):
    n = 60_000
    slab, off, ts = _fuzz_stream(f64, 64, n, SEED + 77 + int(f64))
    schema = bh.telemetry_progress(L.K_DOUBLE if f64 else L.K_INT32)
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, window_ms=500, table_capacity=1 << 15,
                       generic_kernels=generic) as g:
        g.push_batch(slab, off, ts)
        g.flush()
        got = g.snapshot_all()
        k = g.counters(check=False)
    a = O.Agg.run_threads(schema.with_roles(key=KEY), ALL_OPS, 500, SEED, slab, off, ts, 4)
    assert a.hash_collisions() == 0
    assert_records_equal(got, a.snapshot(), ALL_OPS)
    d = a.dropped()
    assert (k["dropped_decode"], k["dropped_enum"], k["dropped_nan"]) == (d["decode"], d["enum"], d["nan"])


@pytest.mark.parametrize("generic", [False, True])
def test_reference_mode_counts_on_mutated_stream(generic):
    n = 60_000
    slab, off, ts = _fuzz_stream(False, 64, n, SEED + 5)
    schema = bh.telemetry_progress()
    with bh.Aggregator(schema, mode=L.MODE_REFERENCE, generic_kernels=generic) as g:
        g.push_batch(slab, off)
        g.flush()
        counts = g.status_counts()
        k = g.counters()
    ref = O.Handler(schema)
    ref.handle_slab(slab, off)
    assert counts == ref.counts()
    assert {"decode": k["dropped_decode"], "enum": k["dropped_enum"]} == ref.dropped()


def _grid_message(rng, total):
    """a message of exactly `total` bytes around the layout the decoder's block tier takes (string key, optional
    unknown pad field, one-byte status and value, optional trailing string), with a near-miss now and then"""
    klen = int(rng.choice([36, 36, 36, 8, 20, 0, 35]))
    key = b"\x0a" + bytes([klen]) + bytes(rng.integers(0x30, 0x7A, klen, dtype=np.uint8))
    status = b"\x10" + bytes([int(rng.integers(0, 8))])
    value = b"\x18" + bytes([int(rng.integers(0, 101))])
    host = b""
    if rng.random() < 0.7:
        hl = int(rng.integers(0, 12))
        host = b"\x22" + bytes([hl]) + bytes(rng.integers(0x61, 0x7A, hl, dtype=np.uint8))
    body = len(key) + len(status) + len(value) + len(host)
    pad = b""
    if total - body >= 2:
        pl = total - body - 2
        pad = b"\x7a" + bytes([pl]) + b"." * pl
    order = rng.random()
    if order < 0.8:
        msg = key + pad + status + value + host
    elif order < 0.9:
        msg = key + status + value + host + pad            # pad at the end: the trailing-field rule
    else:
        msg = status + key + pad + value + host            # another field order: not for the block tier
    m = bytearray(msg)
    r = rng.random()
    if r < 0.25 and m:                                     # near misses: one byte changed somewhere that matters
        i = int(rng.choice([0, 1, 2 + klen, 3 + klen, len(m) - 1, len(m) - 2, int(rng.integers(0, len(m)))]))
        i = min(max(i, 0), len(m) - 1)
        m[i] = int(rng.choice([0x80, 0xFF, 0x00, 0x10, 0x18, 0x22, 0x7A, m[i] ^ 0x80, m[i] + 1 & 0xFF]))
    elif r < 0.30:
        m[-1:] = b""                                       # one byte short: the trailing string now clamps
        m += b"\x00"
    return bytes(m)


@pytest.mark.parametrize("mode", ["windowed", "reference"])
def test_block_tier_messages_on_a_four_byte_grid(mode):
    """sizes that are multiples of 4 keep every message 4-byte aligned in the warp buffer; whether it also stays
    inside one 64-byte block varies from message to message, so the block tier and the tiers below it share
    the batch.  Event by event against the oracle, then through the fused kernel."""
    import torch
    rng = np.random.default_rng(SEED + 4242)
    n = 40_000
    msgs = [_grid_message(rng, int(rng.choice([64, 64, 64, 48, 52, 56, 60, 44, 32, 16]))) for _ in range(n)]
    slab, off = pack(msgs)
    ts = (np.arange(n, dtype=np.uint64) // 7) + 1_700_000_000_000
    schema = bh.telemetry_progress()
    p = torch.from_numpy(np.concatenate([slab, np.zeros(64, np.uint8)])).cuda()
    o = torch.from_numpy(off.view(np.int32)).cuda()
    if mode == "windowed":
        kh = torch.empty(n, dtype=torch.int64, device="cuda"); val = torch.empty(n, dtype=torch.float64, device="cuda")
        st = torch.empty(n, dtype=torch.int32, device="cuda"); oc = torch.empty(n, dtype=torch.uint8, device="cuda")
        with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS) as g:
            g.decode_device(p, o, n, kh, val, st, oc)
            g.sync()
        wkh, wval, wst, woc = O.decode_columns(schema.with_roles(key=KEY), ALL_OPS, SEED, slab, off)
        bad = np.nonzero(oc.cpu().numpy() != woc)[0]
        assert bad.size == 0, (bad[:5], [bytes(slab[off[i]:off[i + 1]]).hex() for i in bad[:3]])
        bad = np.nonzero(kh.cpu().numpy().view(np.uint64) != wkh)[0]
        assert bad.size == 0, (bad[:5], [bytes(slab[off[i]:off[i + 1]]).hex() for i in bad[:3]])
        assert np.array_equal(val.cpu().numpy().view(np.uint64), wval.view(np.uint64))
        assert np.array_equal(st.cpu().numpy(), wst)
        assert (woc == 0).sum() > n // 2 and (woc != 0).sum() > n // 50
        for generic in (False, True):
            with bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, window_ms=500, table_capacity=1 << 16,
                               generic_kernels=generic) as g:
                g.push_batch(slab, off, ts)
                g.flush()
                got = g.snapshot_all()
            a = O.Agg.run_threads(schema.with_roles(key=KEY), ALL_OPS, 500, SEED, slab, off, ts, 4)
            assert a.hash_collisions() == 0
            assert_records_equal(got, a.snapshot(), ALL_OPS)
    else:
        for generic in (False, True):
            with bh.Aggregator(schema, mode=L.MODE_REFERENCE, generic_kernels=generic) as g:
                g.push_batch(slab, off)
                g.flush()
                counts, k = g.status_counts(), g.counters()
            ref = O.Handler(schema)
            ref.handle_slab(slab, off)
            assert counts == ref.counts()
            assert {"decode": k["dropped_decode"], "enum": k["dropped_enum"]} == ref.dropped()

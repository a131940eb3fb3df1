"""The oracle against the committed golden vectors (tests/golden/, made by make_golden.py).

kat_decode.json holds hand-declared expectations derived from protobufjs 6.8.8's reader
behaviour (proto.decode, index.js:129) and the handler's drop rules (index.js:134-152)."""
import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_host
from oracle import oracle as O
from helpers import ALL_OPS, SEED, GOLDEN, load_json, stream_digest, typed_schema

KAT = load_json("kat_decode.json")
EXPECT = load_json("stream_expect.json")


def _cases(group):
    return [pytest.param(c, id=c["name"]) for c in KAT[group]]


def _check(schema, c):
    payload = bytes.fromhex(c["payload"])
    h = O.Handler(schema)
    assert h.handle(payload) == c["outcome"]
    assert h.acks() == 1                      # index.js:151,154 — acked whether counted or not
    m = O.decode(O.schema_c(schema), payload)
    if c["outcome"] == O.EV_DROP_DECODE:
        assert m.outcome == O.EV_DROP_DECODE
        assert h.counts() == {}
        return
    assert m.outcome == O.EV_OK and m.status == c["status"]
    if c["value"] is None:
        assert m.value != m.value
    else:
        assert m.value == c["value"]
    assert payload[m.part_pos[0]: m.part_pos[0] + m.part_len[0]].hex() == c["media_id"]
    name = schema.enum_name(c["status"])
    if c["outcome"] == O.EV_OK:
        assert h.counts() == {name.lower(): 1}  # index.js:136-138
    else:
        assert name is None and h.counts() == {}


@pytest.mark.parametrize("c", _cases("default"))
def test_kat_default_schema(c):
    _check(bh.telemetry_progress().with_roles(key=("mediaId",)), c)


@pytest.mark.parametrize("c", _cases("typed"))
def test_kat_typed_schema(c):
    _check(typed_schema(), c)


def _spec(d):
    return SynthSpec(**d)


@pytest.mark.parametrize("tag", ["c1", "small", "f64_zipf"])
def test_streams_match_golden(tag):
    e = EXPECT[tag]
    spec = _spec(e["spec"])
    slab, off, ts = generate_host(spec, 0, e["n"])
    assert stream_digest(slab, off, ts) == e["sha256"]          # generator determinism
    schema = bh.telemetry_progress(L.K_DOUBLE if spec.value_f64 else L.K_INT32)
    h = O.Handler(schema)
    h.handle_slab(slab, off)
    assert h.counts() == e["handler_counts"]
    assert h.dropped() == e["handler_dropped"]
    assert h.acks() == e["n"]
    assert sum(h.counts().values()) + sum(h.dropped().values()) == e["n"]
    agg = O.Agg(schema.with_roles(key=("mediaId", "status")), ALL_OPS | L.OP_HIST, 10_000, SEED)
    agg.push_raw(slab, off, ts)
    recs, hist = agg.snapshot(with_hist=True)
    import hashlib
    assert len(recs) == e["agg_rows"]
    assert int(recs["count"].sum()) == e["agg_count_total"]
    assert hashlib.sha256(recs[["key_hash", "window_id", "count", "min", "max"]].tobytes()).hexdigest() == e["agg_records_sha256"]
    assert hashlib.sha256(hist.tobytes()).hexdigest() == e["agg_hist_sha256"]
    assert agg.dropped() == e["agg_dropped"]


def test_stored_small_stream():
    z = np.load(f"{GOLDEN}/stream_small.npz")
    schema = bh.telemetry_progress().with_roles(key=("mediaId", "status"))
    agg = O.Agg(schema, ALL_OPS | L.OP_HIST, 10_000, SEED)
    agg.push_raw(z["slab"], z["offsets"], z["ts"])
    recs, hist = agg.snapshot(with_hist=True)
    assert np.array_equal(recs, z["records"].view(recs.dtype)) or all(
        np.array_equal(recs[f], z["records"][f]) for f in ("key_hash", "window_id", "count", "sum", "min", "max"))
    assert np.array_equal(hist, z["hist"])

"""Generates tests/golden/*.json|npz.  Run from the repo root:  python tests/golden/make_golden.py

Two kinds of vectors:
  * kat_decode.json — hand-assembled protobuf byte strings whose expected outcome is DECLARED HERE
    by hand from protobufjs 6.8.8's published reader behaviour (not computed by the oracle); the
    script refuses to write the file if the oracle disagrees with a declaration.
  * stream_*.npz / stream_expect.json — seeded synthetic streams (bh_synth_event_host) with the
    oracle's results at the commit that generated them (regression pins; the reference itself has
    no fixtures to compare with — "parity unpinned").
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import beholder_b200 as bh  # noqa: E402
from beholder_b200 import _lib as L  # noqa: E402
from beholder_b200.synth import SynthSpec, generate_host  # noqa: E402
from oracle import oracle as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
OK, DEC, ENUM = 0, 1, 2


def h(s: str) -> bytes:
    return bytes.fromhex(s.replace(" ", ""))


# (name, payload, handler outcome, status, value, mediaId bytes)   -- default assumed schema
KATS = [
    ("valid_all_fields", h("0a03616263 1004 1832 22026831"), OK, 4, 50.0, b"abc"),
    ("empty_payload_is_all_defaults", b"", OK, 0, 0.0, b""),
    ("missing_status_defaults_to_0", h("0a0178 1807"), OK, 0, 7.0, b"x"),
    ("unknown_enum_number_dropped", h("1063"), ENUM, 99, 0.0, b""),
    ("negative_enum_10_byte_varint", h("10ffffffffffffffffff01"), ENUM, -1, 0.0, b""),
    ("truncated_long_varint_throws", h("10ffffffffffff"), DEC, 0, 0.0, b""),
    ("truncated_string_is_clamped_on_node_buffers", h("0a0a616263"), OK, 0, 0.0, b"abc"),
    ("unknown_varint_field_skipped", h("289601 1002"), OK, 2, 0.0, b""),
    ("unknown_fixed64_field_skipped", h("290102030405060708 1001"), OK, 1, 0.0, b""),
    ("unknown_fixed32_truncated_throws", h("2d0102"), DEC, 0, 0.0, b""),
    ("unknown_len_field_skipped", h("320378797a 1003"), OK, 3, 0.0, b""),
    ("unknown_len_field_truncated_throws", h("32057879"), DEC, 0, 0.0, b""),
    ("wire_type_6_throws", h("36"), DEC, 0, 0.0, b""),
    ("wire_type_7_throws", h("3f"), DEC, 0, 0.0, b""),
    ("stray_end_group_throws", h("34"), DEC, 0, 0.0, b""),
    ("group_skipped_until_any_end_group", h("33 0801 3c 1005"), OK, 5, 0.0, b""),
    ("nested_groups_skipped", h("33 33 34 34 1001"), OK, 1, 0.0, b""),
    ("unterminated_group_throws", h("33 0801"), DEC, 0, 0.0, b""),
    ("known_field_read_by_declared_type_not_wire_type", h("0d026162"), OK, 0, 0.0, b"ab"),
    ("status_tag_with_len_wire_type_still_read_as_varint", h("1203"), OK, 3, 0.0, b""),
    ("repeated_scalar_last_wins", h("1001 1004"), OK, 4, 0.0, b""),
    ("field_number_0_is_just_unknown", h("0005"), OK, 0, 0.0, b""),
    ("non_minimal_varint", h("108400"), OK, 4, 0.0, b""),
    ("negative_progress", h("18fbffffffffffffffff01"), OK, 0, -5.0, b""),
    ("five_byte_varint_all_ones", h("18ffffffff0f"), OK, 0, -1.0, b""),
    ("fifth_byte_high_bits_ignored", h("18ffffffff7f"), OK, 0, -1.0, b""),
    ("six_byte_varint_needs_ten", h("18808080808001 1002"), DEC, 0, 0.0, b""),
    ("six_byte_varint_swallows_five_bytes", h("18808080808001 1002 0000"), OK, 0, 0.0, b""),
    ("varint_skip_unknown_runs_past_end_throws", h("28ffff"), DEC, 0, 0.0, b""),
    ("pad_field_15_skipped", h("0a0161 7a042e2e2e2e 1002 1864"), OK, 2, 100.0, b"a"),
    ("deployed", h("1004"), OK, 4, 0.0, b""),
    ("errored_is_last_known", h("1005"), OK, 5, 0.0, b""),
    ("first_unknown_after_enum_end", h("1006"), ENUM, 6, 0.0, b""),
]

# typed-field schema: progress declared double; plus a bytes field 5 and sint32 field 6, float 7, bool 8, uint32 9
TYPED_FIELDS = [("mediaId", 1, L.K_STRING, L.R_KEY), ("status", 2, L.K_INT32, L.R_STATUS),
                ("progress", 3, L.K_DOUBLE, L.R_VALUE), ("host", 4, L.K_STRING, 0), ("blob", 5, L.K_BYTES, 0),
                ("delta", 6, L.K_SINT32, 0), ("ratio", 7, L.K_FLOAT, 0), ("flag", 8, L.K_BOOL, 0), ("n", 9, L.K_UINT32, 0)]
KATS_TYPED = [
    ("double_value", h("19 000000000000f83f"), OK, 0, 1.5, b""),
    ("double_truncated_throws", h("19 0000000000"), DEC, 0, 0.0, b""),
    ("double_nan_decodes", h("19 000000000000f87f"), OK, 0, float("nan"), b""),
    ("bytes_field_strict_truncation_throws", h("2a05 6162"), DEC, 0, 0.0, b""),
    ("bytes_field_ok", h("2a02 6162 1002"), OK, 2, 0.0, b""),
    ("sint32_float_bool_uint32_read", h("3003 3d0000803f 4001 48ffffffff0f 1003"), OK, 3, 0.0, b""),
    ("float_truncated_throws", h("3d0000"), DEC, 0, 0.0, b""),
]


def typed_schema():
    return bh.Schema([bh.Field(*f) for f in TYPED_FIELDS], dict(bh.telemetry_progress().enum))


def check_kats(schema, kats):
    sc = O.schema_c(schema)
    out = []
    for name, payload, outcome, status, value, media in kats:
        m = O.decode(sc, payload)
        hd = O.Handler(schema)
        got = hd.handle(payload)
        assert got == outcome, (name, got, outcome)
        assert hd.acks() == 1, name
        if outcome != DEC:
            assert m.outcome == 0 and m.status == status, (name, m.outcome, m.status)
            assert (m.value == value) or (value != value and m.value != m.value), (name, m.value)
            if m.n_parts:
                assert payload[m.part_pos[0]: m.part_pos[0] + m.part_len[0]] == media, name
        else:
            assert m.outcome == 1, name
        out.append({"name": name, "payload": payload.hex(), "outcome": outcome, "status": status,
                    "value": None if value != value else value, "media_id": media.hex()})
    return out


def stream_expect(tag, spec, n, store_stream):
    slab, off, ts = generate_host(spec, 0, n)
    digest = hashlib.sha256(slab[: off[-1]].tobytes() + off.tobytes() + ts.tobytes()).hexdigest()
    schema = bh.telemetry_progress(L.K_DOUBLE if spec.value_f64 else L.K_INT32)
    hd = O.Handler(schema)
    hd.handle_slab(slab, off)
    ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX | L.OP_HIST
    agg = O.Agg(schema.with_roles(key=("mediaId", "status")), ops, 10_000, 0xBE401DE5)
    agg.push_raw(slab, off, ts)
    recs, hist = agg.snapshot(with_hist=True)
    assert agg.hash_collisions() == 0
    if store_stream:
        np.savez_compressed(os.path.join(HERE, f"stream_{tag}.npz"), slab=slab, offsets=off, ts=ts, records=recs, hist=hist)
    return {
        "spec": spec.__dict__, "n": n, "sha256": digest,
        "handler_counts": hd.counts(), "handler_dropped": {k: int(v) for k, v in hd.dropped().items()},
        "agg_rows": int(len(recs)), "agg_dropped": {k: int(v) for k, v in agg.dropped().items()},
        "agg_count_total": int(recs["count"].sum()),
        "agg_records_sha256": hashlib.sha256(recs[["key_hash", "window_id", "count", "min", "max"]].tobytes()).hexdigest(),
        "agg_hist_sha256": hashlib.sha256(hist.tobytes()).hexdigest(),
        "agg_sum_total": float(recs["sum"].sum()),
    }


def main():
    kat = {"schema": "assumed api.TelemetryProgress (beholder_b200.schema.telemetry_progress)",
           "default": check_kats(bh.telemetry_progress().with_roles(key=("mediaId",)), KATS),
           "typed": check_kats(typed_schema(), KATS_TYPED)}
    with open(os.path.join(HERE, "kat_decode.json"), "w") as f:
        json.dump(kat, f, indent=1)
    expect = {
        # BASELINE.json config 1: 10k events, reference schema, 1 % malformed + 1 % unknown enum
        "c1": stream_expect("c1", SynthSpec(n_keys=64, rate_per_ms=1, malformed_ppm=10000, unknown_enum_ppm=10000), 10000, False),
        # a small stream stored in full (independent of the generator staying unchanged)
        "small": stream_expect("small", SynthSpec(n_keys=40, rate_per_ms=1, payload_bytes=0, malformed_ppm=30000,
                                                 unknown_enum_ppm=30000), 512, True),
        "f64_zipf": stream_expect("f64", SynthSpec(n_keys=200, rate_per_ms=1, value_f64=True, value_signed=True, zipf=True,
                                                   payload_bytes=96), 4096, False),
    }
    with open(os.path.join(HERE, "stream_expect.json"), "w") as f:
        json.dump(expect, f, indent=1)
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()

"""Independent cross-check of the oracle's wire decoder: Python google.protobuf parses the same
well-formed messages (generated streams and hand-built ones) to the same field values.
Malformed inputs are NOT compared: protobufjs 6.8.8 and google.protobuf differ there by design
(known fields read by declared type, lenient string clamp) and the oracle follows protobufjs."""
import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_host
from oracle import oracle as O

pb = pytest.importorskip("google.protobuf")


def _message_class(double_progress=False):
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    fdp = descriptor_pb2.FileDescriptorProto(name=f"t{int(double_progress)}.proto", package="api", syntax="proto3")
    e = fdp.enum_type.add(name="TelemetryStatusEntry")
    for n, v in bh.telemetry_progress().enum.items():
        e.value.add(name=n, number=v)
    m = fdp.message_type.add(name="TelemetryProgress")
    m.field.add(name="mediaId", number=1, type=9, label=1)
    m.field.add(name="status", number=2, type=14, type_name=".api.TelemetryStatusEntry", label=1)
    m.field.add(name="progress", number=3, type=1 if double_progress else 5, label=1)
    m.field.add(name="host", number=4, type=9, label=1)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("api.TelemetryProgress"))


@pytest.mark.parametrize("f64", [False, True])
@pytest.mark.parametrize("payload_bytes", [0, 64, 300])
def test_generated_events_parse_identically(f64, payload_bytes):
    cls = _message_class(f64)
    spec = SynthSpec(n_keys=77, rate_per_ms=3, payload_bytes=payload_bytes, value_f64=f64, value_signed=True,
                     unknown_enum_ppm=20000)
    slab, off, ts = generate_host(spec, 1000, 1500)
    schema = bh.telemetry_progress(L.K_DOUBLE if f64 else L.K_INT32).with_roles(key=("mediaId", "host"))
    sc = O.schema_c(schema)
    for i in range(1500):
        b = bytes(slab[off[i]: off[i + 1]])
        if payload_bytes:   # padded to exactly payload_bytes unless the natural size is within 1 byte of it or larger
            assert len(b) == payload_bytes or len(b) >= payload_bytes - 1
        g = cls.FromString(b)
        m = O.decode(sc, b)
        assert m.outcome == O.EV_OK
        assert m.status == g.status
        assert m.value == g.progress
        assert b[m.part_pos[0]: m.part_pos[0] + m.part_len[0]] == g.mediaId.encode()
        assert b[m.part_pos[1]: m.part_pos[1] + m.part_len[1]] == g.host.encode()


def test_python_encoder_round_trip():
    cls = _message_class(False)
    schema = bh.telemetry_progress()
    sc = O.schema_c(schema.with_roles(key=("mediaId",)))
    rng = np.random.default_rng(7)
    for _ in range(300):
        g = cls(mediaId="m%d" % rng.integers(1 << 40), status=int(rng.integers(0, 6)),
                progress=int(rng.integers(-5, 200)), host="h" * int(rng.integers(0, 5)))
        b = g.SerializeToString()
        m = O.decode(sc, b)
        assert (m.outcome, m.status, m.value) == (0, g.status, g.progress)
        b2 = bh.encode(schema, mediaId=g.mediaId, status=g.status, progress=g.progress, host=g.host)
        g2 = cls.FromString(b2)
        assert g2 == g

"""AMQP 0-9-1 delivery tap: bh_amqp_feed against the Python restatement in
oracle/amqp_oracle.py on hand-built broker streams.  Host code only -- runs without a GPU."""
import random
import struct

import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import amqp as A
from beholder_b200 import _lib as L
from oracle import amqp_oracle as AO
from helpers import SEED


def _collect(tap, pieces):
    got = []
    for piece in pieces:
        for slab, off, d in tap.feed(piece):
            assert off[0] == 0 and np.all(np.diff(off.astype(np.int64)) == d["body_bytes"])
            for i in range(len(d)):
                got.append((int(d["delivery_tag"][i]), int(d["channel"][i]), int(d["redelivered"][i]),
                            bytes(slab[off[i]:off[i + 1]])))
    return got


def _want(stream, consumer_tag=None):
    return [(w.delivery_tag, w.channel, w.redelivered, w.body) for w in AO.scan(stream, consumer_tag)]


def _split(stream, rng, max_piece):
    out, i = [], 0
    while i < len(stream):
        n = rng.randint(1, max_piece)
        out.append(stream[i:i + n]); i += n
    return out


def _progress_bodies(n, rng):
    schema = bh.telemetry_progress()
    out = []
    for i in range(n):
        out.append(bh.encode(schema, mediaId="%036d" % rng.randrange(50), status=rng.randrange(6),
                             progress=rng.randrange(101), host="h%d" % rng.randrange(9)))
    return out


def test_whole_stream_and_byte_at_a_time():
    rng = random.Random(SEED)
    bodies = _progress_bodies(200, rng) + [b"", b"x" * 9000, b"y" * 4088, b"z" * 4089]
    stream = A.encode_stream(bodies, heartbeat_every=7, frame_max=4096)
    want = _want(stream)
    assert len(want) == len(bodies) and [w[3] for w in want] == bodies
    with A.AmqpTap() as tap:
        assert _collect(tap, [stream]) == want
        s = tap.stats()
        assert s["deliveries"] == len(bodies) and s["heartbeats"] == (len(bodies) + 6) // 7
        assert s["body_bytes"] == sum(map(len, bodies))
    with A.AmqpTap() as tap:
        assert _collect(tap, [stream[i:i + 1] for i in range(len(stream))]) == want


@pytest.mark.parametrize("max_piece", [3, 64, 1500, 70000])
def test_random_socket_reads(max_piece):
    rng = random.Random(SEED + max_piece)
    bodies = _progress_bodies(500, rng)
    for i in range(0, 500, 50):
        bodies[i] = bytes(rng.randrange(256) for _ in range(rng.randrange(20000)))   # multi-frame bodies
    stream = A.encode_stream(bodies, heartbeat_every=11, frame_max=4096, redelivered=True)
    with A.AmqpTap() as tap:
        assert _collect(tap, _split(stream, rng, max_piece)) == _want(stream)


def test_channels_interleaved_and_consumer_filter():
    rng = random.Random(SEED)
    bodies = _progress_bodies(60, rng)
    big = b"B" * 10000
    # channel 1: progress consumer; channel 2: another consumer; a large message on channel 2 whose body
    # frames are interleaved with whole deliveries on channel 1 (frames of different channels may mix)
    parts = []
    big_frames = A.encode_delivery(big, 900, channel=2, consumer_tag="status-consumer", frame_max=4096)
    cut = []
    i = 0
    while i < len(big_frames):        # split into frames
        size = struct.unpack_from(">I", big_frames, i + 3)[0]
        cut.append(big_frames[i:i + 8 + size]); i += 8 + size
    for k, b in enumerate(bodies):
        parts.append(A.encode_delivery(b, k + 1, channel=1, consumer_tag="progress-consumer"))
        if k < len(cut):
            parts.append(cut[k])
    parts += cut[len(bodies):]
    stream = b"".join(parts)
    want_all = _want(stream)
    assert len(want_all) == len(bodies) + 1
    with A.AmqpTap() as tap:
        assert sorted(_collect(tap, _split(stream, rng, 700))) == sorted(want_all)
    with A.AmqpTap("progress-consumer") as tap:
        got = _collect(tap, _split(stream, rng, 700))
        assert got == _want(stream, "progress-consumer") and len(got) == len(bodies)
        assert tap.stats()["filtered"] == 1


def test_other_methods_and_returned_content_are_skipped():
    body = b"payload"
    stream = (A.encode_method(0, 10, 60, b"\x00")                      # connection.blocked
              + A.encode_method(1, 60, 50, struct.pack(">H", 312) + b"\x08NO_ROUTE" + b"\x02ex" + b"\x02rk")   # basic.return
              + A.encode_content(1, b"returned message")
              + A.encode_method(1, 60, 71, b"\x00" * 12) + A.encode_content(1, b"")   # get-ok, empty body
              + A.encode_delivery(body, 5)
              + A.encode_method(1, 60, 30, b"\x04ctag"))                # basic.cancel
    with A.AmqpTap() as tap:
        got = _collect(tap, [stream])
        assert got == [(5, 1, 0, body)] == _want(stream)
        s = tap.stats()
        assert s["other_methods"] == 2 and s["skipped_content"] == 2 and s["deliveries"] == 1


def test_small_slab_and_message_cap_stop_and_resume():
    rng = random.Random(SEED)
    bodies = _progress_bodies(300, rng)
    stream = A.encode_stream(bodies)
    with A.AmqpTap(slab_bytes=1000, max_messages=7) as tap:
        batches = 0
        got = []
        for slab, off, d in tap.feed(stream):
            assert len(d) <= 7 and int(off[-1]) <= 1000
            batches += 1
            got += [bytes(slab[off[i]:off[i + 1]]) for i in range(len(d))]
        assert got == bodies and batches >= 300 // 7


def test_protocol_errors():
    ok = A.encode_delivery(b"abc", 1)
    bad_end = bytearray(A.encode_delivery(b"abc", 2)); bad_end[-1] = 0x00
    cases = {
        "bad frame end": ok + bytes(bad_end),
        "unknown frame type": ok + A.frame(5, 1, b""),
        "content header on channel 1 without a method": ok + A.encode_content(1, b"abc"),
        "without a header": ok + A.frame(A.FRAME_BODY, 1, b"abc"),
        "exceeds the announced": A.encode_method(1, 60, 60, b"\x01c" + struct.pack(">QB", 3, 0) + b"\x00\x00")
                                 + A.frame(A.FRAME_HEADER, 1, struct.pack(">HHQ", 60, 0, 2) + b"\x00\x00") + A.frame(A.FRAME_BODY, 1, b"abc"),
        "while content is expected": A.encode_method(1, 60, 60, b"\x01c" + struct.pack(">QB", 3, 0) + b"\x00\x00") + A.encode_method(1, 60, 30, b""),
        "truncated basic.deliver": A.encode_method(1, 60, 60, b"\x05c"),
    }
    for text, stream in cases.items():
        with pytest.raises(AO.ProtocolError):
            AO.scan(stream)
        with A.AmqpTap() as tap:
            with pytest.raises(A.AmqpProtocolError) as e:
                _collect(tap, [stream])
            assert text in str(e.value), (text, str(e.value))
            with pytest.raises(A.AmqpProtocolError):      # the tap stays dead
                _collect(tap, [ok])
    with A.AmqpTap(max_body=100) as tap:
        with pytest.raises(A.AmqpProtocolError) as e:
            _collect(tap, [A.encode_delivery(b"x" * 101, 1)])
        assert "max_body" in str(e.value)
    with A.AmqpTap(slab_bytes=50) as tap:
        with pytest.raises(A.AmqpProtocolError) as e:
            _collect(tap, [A.encode_delivery(b"x" * 51, 1)])
        assert e.value.code == L.BH_ERR_TOO_LARGE


@pytest.mark.gpu
def test_tap_feeds_the_gpu_handler_path():
    """broker bytes -> tap -> push_batch -> per-status counts = the oracle handler on the same bodies
    (index.js:127-138), with malformed and unknown-enum bodies in the stream"""
    from beholder_b200.synth import SynthSpec, generate_host
    from oracle import oracle as O
    from helpers import pack
    spec = SynthSpec(n_keys=300, payload_bytes=0, malformed_ppm=20_000, unknown_enum_ppm=20_000)
    n = 20_000
    slab, off, ts = generate_host(spec, 0, n)
    bodies = [bytes(slab[off[i]:off[i + 1]]) for i in range(n)]
    stream = A.encode_stream(bodies, heartbeat_every=500, consumer_tag="progress")
    rng = random.Random(SEED)
    schema = bh.telemetry_progress()
    tags = []
    with bh.Aggregator(schema, mode=L.MODE_REFERENCE) as g, A.AmqpTap("progress", pinned=True) as tap:
        for piece in _split(stream, rng, 200_000):
            for s, o, d in tap.feed(piece):
                g.push_batch(s, o, None, pinned=True)
                g.flush()                       # the pinned slab is reused by the next batch
                tags += [int(t) for t in d["delivery_tag"]]
        counts = g.status_counts()
        k = g.counters()
    ref = O.Handler(schema)
    ref.handle_slab(slab, off)
    assert tags == list(range(1, n + 1))
    assert counts == ref.counts()
    assert {"decode": k["dropped_decode"], "enum": k["dropped_enum"]} == ref.dropped()

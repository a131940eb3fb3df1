"""AMQP 0-9-1 delivery tap: raw broker -> client bytes in, (slab, offsets, delivery tags) out.

Stands where amqplib's frame parser and `amqp.listen(topic, fn)` (index.js:43-44,127) stand in the
reference: instead of one `rmsg` object per delivery, the bodies of a whole socket read land in the
slab/offsets form `Aggregator.push_batch` takes, and the delivery tags come back for the batched
`ack` (index.js:154) once the batch is committed.  The scan itself is `bh_amqp_feed` in
csrc/bh_amqp.cpp; this file wraps it and holds the frame ENCODER used by tests and benchmarks to
build broker-side byte streams.
"""
from __future__ import annotations

import ctypes as C
import struct
from typing import Iterable, List, Optional, Tuple

import numpy as np

from . import _lib as L

FRAME_METHOD, FRAME_HEADER, FRAME_BODY, FRAME_HEARTBEAT, FRAME_END = 1, 2, 3, 8, 0xCE
DELIVERY_DTYPE = np.dtype([("delivery_tag", "<u8"), ("body_bytes", "<u4"), ("channel", "<u2"), ("redelivered", "<u2")])


class AmqpProtocolError(L.BeholderError):
    pass


class AmqpTap:
    """One broker connection's delivery scanner.  `feed(data)` accepts any split of the byte stream
    (socket reads): an incomplete frame at the end is kept and prepended to the next call."""

    def __init__(self, consumer_tag: Optional[str] = None, *, slab_bytes: int = 4 << 20, max_messages: int = 1 << 16,
                 max_body: int = 0, pinned: bool = False):
        self._h = C.c_void_p()
        L.check(L.lib.bh_amqp_create(C.byref(self._h), consumer_tag.encode() if consumer_tag else None, max_body))
        self._pin = None
        if pinned:      # bodies land in page-locked memory: push_batch(..., pinned=True) copies straight from it
            from .engine import PinnedBuffer
            self._pin = PinnedBuffer(slab_bytes)
            self.slab = self._pin.view(np.uint8)
        else:
            self.slab = np.empty(slab_bytes, dtype=np.uint8)
        self.offsets = np.zeros(max_messages + 1, dtype=np.uint32)
        self.deliveries = np.zeros(max_messages, dtype=DELIVERY_DTYPE)
        self._rest = b""

    def close(self) -> None:
        if self._h:
            L.lib.bh_amqp_destroy(self._h)
            self._h = C.c_void_p()
        if self._pin is not None:
            self._pin.free()
            self._pin = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def feed(self, data: bytes) -> Iterable[Tuple[np.ndarray, np.ndarray, np.ndarray]]:
        """Yields (slab, offsets[0..n], deliveries[0..n)) batches; the arrays are views that the next
        iteration overwrites, so consume (push_batch) each batch before asking for the next."""
        buf = self._rest + bytes(data)
        arr = np.frombuffer(buf, dtype=np.uint8) if buf else np.empty(0, dtype=np.uint8)
        pos = 0
        while True:
            consumed, n = C.c_uint64(), C.c_uint32()
            rc = L.lib.bh_amqp_feed(self._h, arr.ctypes.data + pos if len(buf) else None, len(buf) - pos,
                                    self.slab.ctypes.data, self.slab.nbytes, self.offsets.ctypes.data,
                                    self.deliveries.ctypes.data_as(C.POINTER(L.BhAmqpDelivery)), len(self.deliveries),
                                    C.byref(consumed), C.byref(n))
            pos += consumed.value
            if n.value:
                yield self.slab, self.offsets[: n.value + 1], self.deliveries[: n.value]
            if rc != L.BH_OK:
                msg = L.lib.bh_amqp_last_error(self._h).decode()
                self._rest = b""
                raise AmqpProtocolError(rc, msg)
            if n.value == 0 or consumed.value == 0:
                break
        self._rest = buf[pos:]

    def stats(self) -> dict:
        s = L.BhAmqpStats()
        L.check(L.lib.bh_amqp_get_stats(self._h, C.byref(s)))
        return {n: getattr(s, n) for n, _ in s._fields_}


# ---------------------------------------------------------------------------------------------
# broker-side encoder (tests, benchmarks): what RabbitMQ sends for one delivery
# ---------------------------------------------------------------------------------------------
def frame(ftype: int, channel: int, payload: bytes) -> bytes:
    return struct.pack(">BHI", ftype, channel, len(payload)) + payload + bytes([FRAME_END])


def _shortstr(s: str) -> bytes:
    b = s.encode()
    return bytes([len(b)]) + b


def encode_method(channel: int, cls: int, method: int, args: bytes = b"") -> bytes:
    return frame(FRAME_METHOD, channel, struct.pack(">HH", cls, method) + args)


def encode_content(channel: int, body: bytes, frame_max: int = 4096, properties: bytes = b"\x00\x00") -> bytes:
    """content header (class 60, weight 0, body size, property flags) + body frames of <= frame_max - 8"""
    out = [frame(FRAME_HEADER, channel, struct.pack(">HHQ", 60, 0, len(body)) + properties)]
    step = max(1, frame_max - 8)
    for i in range(0, len(body), step):
        out.append(frame(FRAME_BODY, channel, body[i:i + step]))
    return b"".join(out)


def encode_delivery(body: bytes, delivery_tag: int, *, channel: int = 1, consumer_tag: str = "amq.ctag-beholder",
                    exchange: str = "v1.telemetry.progress", routing_key: str = "", redelivered: bool = False,
                    frame_max: int = 4096) -> bytes:
    args = _shortstr(consumer_tag) + struct.pack(">QB", delivery_tag, int(redelivered)) + _shortstr(exchange) + _shortstr(routing_key)
    return encode_method(channel, 60, 60, args) + encode_content(channel, body, frame_max)


def heartbeat() -> bytes:
    return frame(FRAME_HEARTBEAT, 0, b"")


def encode_stream(bodies: List[bytes], *, first_tag: int = 1, heartbeat_every: int = 0, **kw) -> bytes:
    out = []
    for i, b in enumerate(bodies):
        if heartbeat_every and i % heartbeat_every == 0:
            out.append(heartbeat())
        out.append(encode_delivery(b, first_tag + i, **kw))
    return b"".join(out)

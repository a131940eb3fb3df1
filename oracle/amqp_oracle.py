"""CPU restatement of the AMQP 0-9-1 delivery scan (beholder_b200/csrc/bh_amqp.cpp).

TEST INFRASTRUCTURE ONLY -- imported by tests/ alone.  PARITY UNPINNED: the reference gets its
deliveries from amqplib 0.5.5 through triton-core's `amqp.listen` (index.js:43-44,127); neither
package is vendored under /root/reference and the reference has no tests, so this follows the
published AMQP 0-9-1 frame grammar (specification section 4.2; basic.deliver = class 60 method 60)
and what amqplib is documented to do with it: bodies of all body frames of a delivery concatenated
into `message.content`, `fields.deliveryTag` / `fields.redelivered` / `fields.consumerTag` taken from
the method frame, heartbeats dropped, a bad frame end or an unknown frame type fatal.

Written as a pull parser over the complete stream (no incremental state), unlike the C++ tap.
"""
from __future__ import annotations

import struct
from typing import List, NamedTuple, Optional


class Delivery(NamedTuple):
    delivery_tag: int
    channel: int
    redelivered: int
    consumer_tag: str
    body: bytes


class ProtocolError(Exception):
    pass


def _frames(stream: bytes):
    """yields (type, channel, payload) for every complete frame; stops at an incomplete tail"""
    i = 0
    while len(stream) - i >= 8:
        ftype, channel, size = struct.unpack_from(">BHI", stream, i)
        if ftype not in (1, 2, 3, 8):
            raise ProtocolError(f"unknown frame type {ftype}")
        if len(stream) - i < 8 + size:
            return
        if stream[i + 7 + size] != 0xCE:
            raise ProtocolError("bad frame end")
        yield ftype, channel, stream[i + 7:i + 7 + size]
        i += 8 + size


def scan(stream: bytes, consumer_tag: Optional[str] = None) -> List[Delivery]:
    """deliveries completed inside `stream`, in completion order"""
    out: List[Delivery] = []
    pending = {}      # channel -> [kind, fields, size, chunks]
    for ftype, ch, pay in _frames(stream):
        if ftype == 8:
            continue
        if ftype == 1:
            if ch in pending:
                raise ProtocolError("method while content expected")
            cls, mth = struct.unpack_from(">HH", pay, 0)
            if (cls, mth) == (60, 60):
                try:
                    q = 4
                    n = pay[q]; ctag = pay[q + 1:q + 1 + n].decode(); q += 1 + n
                    tag, red = struct.unpack_from(">QB", pay, q); q += 9
                    n = pay[q]; q += 1 + n            # exchange
                    n = pay[q]; q += 1 + n            # routing key
                except (IndexError, struct.error):
                    raise ProtocolError("truncated deliver")
                if q > len(pay):
                    raise ProtocolError("truncated deliver")
                pending[ch] = ["deliver", (tag, red & 1, ctag), None, []]
            elif (cls, mth) in ((60, 50), (60, 71)):
                pending[ch] = ["skip", None, None, []]
        elif ftype == 2:
            if ch not in pending or pending[ch][2] is not None:
                raise ProtocolError("header without method")
            (size,) = struct.unpack_from(">Q", pay, 4)
            pending[ch][2] = size
        else:
            if ch not in pending or pending[ch][2] is None:
                raise ProtocolError("body without header")
            pending[ch][3].append(pay)
            if sum(map(len, pending[ch][3])) > pending[ch][2]:
                raise ProtocolError("body exceeds announced size")
        if ch in pending and pending[ch][2] is not None and sum(map(len, pending[ch][3])) == pending[ch][2]:
            kind, fields, _, chunks = pending.pop(ch)
            if kind == "deliver" and (not consumer_tag or fields[2] == consumer_tag):
                out.append(Delivery(fields[0], ch, fields[1], fields[2], b"".join(chunks)))
    return out

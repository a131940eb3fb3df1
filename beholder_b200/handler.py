"""Host-side mirror of the reference's handler interface for the progress path.

The reference (JavaScript, index.js) wires the path like this:

    const prom = Prom.new('beholder'); Prom.expose()                            // index.js:27-28
    metrics.progress_updates_total = new prom.Counter({name, help, labelNames}) // index.js:30-34
    amqp.listen('v1.telemetry.progress', async rmsg => { ...; rmsg.ack() })     // index.js:127-155

Node is not available in this image, so the host side above the C-ABI is written in Python
with the same names, argument meaning and error behaviour; addon/ holds the N-API shim and the
JS wrapper a maintainer would ship.  The drop-in is:

    gpu = BeholderGPU()                                     # replaces the Counter + decode + inc
    amqp.listen('v1.telemetry.progress', gpu.handler(side_effects))
    gpu.progress_updates_total.get()                        # what the /metrics scrape reads

`handler` returns an `async rmsg => ...` callable: it copies `rmsg.message.content` into the
pinned ingest ring (bh_push), keeps `rmsg` and acks it exactly once when the batch holding it
has been aggregated on the GPU (index.js:151,154: every message is acked, dropped or not).

Ack liveness.  The broker stops delivering once `prefetch` messages are unacknowledged (index.js:43,
`new AMQP(addr, 100, ...)`), so acks must never wait for more traffic.  Three things guarantee that:
the handler submits the partial batch as soon as half a prefetch window is waiting (bh_submit); the
library submits a partial batch whose oldest message is `max_batch_latency_ms` old; and the
library's completion thread reports finished batches through a callback, which is posted to the
event loop (`call_soon_threadsafe`, the counterpart of napi_threadsafe_function) and sends the acks.
Worst-case ack latency = max_batch_latency_ms + one kernel launch (tens of microseconds).
"""
from __future__ import annotations

import asyncio
import collections
import inspect
from typing import Awaitable, Callable, Deque, Dict, List, Optional, Tuple

from . import _lib as L
from .engine import Aggregator
from .schema import Schema, decode_message, telemetry_progress, telemetry_status

COUNTER_NAME = "beholder_progress_updates_total"                                       # index.js:31
COUNTER_HELP = "Total number of messages processed in this processes lifetime"         # index.js:32


class GpuCounter:
    """prom-client Counter look-alike whose values live in the GPU table (index.js:30-34)."""

    def __init__(self, owner: "BeholderGPU", name: str = COUNTER_NAME, help: str = COUNTER_HELP,
                 label_names=("status",), counts=None):
        self._owner = owner
        self._counts = counts or owner.status_counts
        self.name = name
        self.help = help
        self.labelNames = list(label_names)

    def get(self) -> dict:
        """same shape as prom-client 11.x Counter.get()"""
        counts = self._counts()
        return {
            "help": self.help, "name": self.name, "type": "counter", "aggregator": "sum",
            "values": [{"value": v, "labels": {"status": k}} for k, v in counts.items()],
        }

    def inc(self, labels=None, value=1):
        raise TypeError("beholder_progress_updates_total is incremented on the GPU by the progress handler; "
                        "there is no host-side inc()")

    def metrics(self) -> str:
        """prom-client registry.metrics() text for this counter"""
        lines = [f"# HELP {self.name} {self.help}", f"# TYPE {self.name} counter"]
        for k, v in self._counts().items():
            lines.append(f'{self.name}{{status="{k}"}} {v}')
        return "\n".join(lines) + "\n"


class HostCounter:
    """plain host counter for the side-effect metrics (beholder_trello_comments, index.js:35-39)"""

    def __init__(self, name: str, help: str = ""):
        self.name, self.help, self.value = name, help, 0

    def inc(self, value: int = 1):
        self.value += value

    def metrics(self) -> str:
        return f"# HELP {self.name} {self.help}\n# TYPE {self.name} counter\n{self.name} {self.value}\n"


class BeholderGPU:
    """`require('beholder-gpu')(cfg)` of SURVEY §8b."""

    def __init__(self, schema: Optional[Schema] = None, device: int = 0, commit_every: int = 65536,
                 slab_bytes: int = 0, slab_events: int = 0, logger=None, prefetch: int = 100,
                 max_batch_latency_ms: int = 5):
        self.schema = schema or telemetry_progress()
        self.agg = Aggregator(self.schema, mode=L.MODE_REFERENCE, device=device,
                              slab_bytes=slab_bytes, slab_events=slab_events,
                              max_batch_latency_ms=max_batch_latency_ms)
        self.commit_every = commit_every
        self.logger = logger
        self._init_ack_state(prefetch)
        self.progress_updates_total = GpuCounter(self)
        self.status_agg = None
        self.trello_comments_total = HostCounter("beholder_trello_comments",
                                                 "Total trello comments crreated in this processes lifetime")

    def _reaches_side_effects(self, content) -> bool:
        """The reference reaches db.getByID and the Trello comment (index.js:140-148) only after proto.decode
        (:129) and enumToString(...).toLowerCase() (:134,137) succeeded; a message that does not decode, or whose
        status has no enum name, goes to the catch (:149) with no side effect.  Same gate here, decided on the host
        for this one message (the GPU decodes the same bytes for the counters)."""
        msg = decode_message(self.schema, bytes(content))
        return msg is not None and self.schema.enum_name(msg.get("status", 0)) is not None

    def _init_ack_state(self, prefetch: int):
        self._pending: Deque[Tuple[int, object]] = collections.deque()   # (sequence number, rmsg)
        self._seq = 0
        self._unsubmitted = 0
        # hand a batch to the GPU once half of the broker's prefetch window (index.js:43) waits for its ack
        self.submit_at = max(1, min(self.commit_every, prefetch // 2 if prefetch else self.commit_every))
        self._loop = None

    def _attach_loop(self):
        """first delivery: route the library's completion callback onto this event loop"""
        self._loop = loop = asyncio.get_running_loop()

        def on_commit(seq):                            # runs on the library's completion thread
            try:
                loop.call_soon_threadsafe(self._ack_upto, seq)
            except RuntimeError:                       # loop already closed: commit() / close() will ack
                pass
        self.agg.set_commit_callback(on_commit)

    # -- the handler: amqp.listen('v1.telemetry.progress', gpu.handler(fn)) --
    def handler(self, side_effects: Optional[Callable[[object], Optional[Awaitable]]] = None):
        async def on_progress(rmsg):
            if self._loop is None:
                self._attach_loop()
            content = rmsg.message.content
            self.agg.push(bytes(content))
            self._seq += 1
            self._unsubmitted += 1
            self._pending.append((self._seq, rmsg))
            if side_effects is not None and self._reaches_side_effects(content):
                # index.js:140-148 stay on the host, per event; their failure neither un-counts the
                # event nor stops the ack (index.js:149-152)
                try:
                    r = side_effects(rmsg)
                    if inspect.isawaitable(r):
                        await r
                except Exception as err:  # noqa: BLE001 - mirrors `catch (err)` at index.js:149
                    if self.logger:
                        self.logger.warning("failed to update media progress %s", err)
            if len(self._pending) >= self.commit_every:
                self.commit()
                return
            if self._unsubmitted >= self.submit_at:
                self._unsubmitted = 0
                self.agg.submit()
            self._ack_upto(self.agg.poll_committed())
        return on_progress

    # -- v1.telemetry.status (index.js:62-125): per-status counting of the sibling stream.
    # The reference does NOT count this stream; `beholder_status_updates_total` is an added series.  Its
    # I/O (db.updateStatus, Trello move, Telegram, Emby) stays on the host in `side_effects`; as in the
    # reference that code is outside any try/catch (index.js:63-68), so an exception propagates and the
    # message is not acked.  Unlike the reference, a message whose payload cannot be decoded is counted as
    # dropped and acked instead of being redelivered forever.
    def status_handler(self, side_effects: Optional[Callable[[object], Optional[Awaitable]]] = None,
                       schema: Optional[Schema] = None):
        if self.status_agg is None:
            self.status_schema = schema or telemetry_status()
            self.status_agg = Aggregator(self.status_schema, mode=L.MODE_REFERENCE, device=self.agg.device)
            self.status_updates_total = GpuCounter(self, "beholder_status_updates_total",
                                                   "Total number of status messages processed in this processes lifetime",
                                                   counts=self.status_counts_sibling)

        async def on_status(rmsg):
            if side_effects is not None:
                r = side_effects(rmsg)
                if inspect.isawaitable(r):
                    await r
            self.status_agg.push(bytes(rmsg.message.content))
            rmsg.ack()                                    # index.js:124
        return on_status

    def status_counts_sibling(self) -> Dict[str, int]:
        self.status_agg.flush()
        out = {}
        for r in self.status_agg.snapshot(0):
            name = self.status_schema.enum_name(int(r["key_hash"]))
            if name is not None:
                out[name.lower()] = int(r["count"])
        return out

    def _ack_upto(self, seq: int) -> int:
        n = 0
        while self._pending and self._pending[0][0] <= seq:
            _, rmsg = self._pending.popleft()
            rmsg.ack()
            n += 1
        return n

    def commit(self) -> int:
        """aggregate everything received so far, then ack it; returns the number of acks sent"""
        self._unsubmitted = 0
        return self._ack_upto(self.agg.flush())

    def status_counts(self) -> Dict[str, int]:
        self.commit()
        return self.agg.status_counts()

    def dropped(self) -> Dict[str, int]:
        k = self.agg.counters()
        return {"decode": k["dropped_decode"], "enum": k["dropped_enum"]}

    def metrics(self) -> str:
        """the body Prom.expose() serves for these two counters (index.js:28)"""
        return self.progress_updates_total.metrics() + self.trello_comments_total.metrics()

    def close(self):
        if self._loop is not None:
            self.agg.set_commit_callback(None)         # no more calls into a loop that may be gone
        self.commit()
        self.agg.close()
        if self.status_agg is not None:
            self.status_agg.close()


# --------------------------------------------------------------------------------------------
# local stub queue: stands in for triton-core/amqp in tests and BASELINE.json config 1
# --------------------------------------------------------------------------------------------
class _Message:
    __slots__ = ("content",)

    def __init__(self, content: bytes):
        self.content = content


class StubRmsg:
    """the `rmsg` triton-core/amqp hands to a listener: .message.content and .ack()"""
    __slots__ = ("message", "acks", "_queue")

    def __init__(self, content: bytes, queue: "StubAMQP"):
        self.message = _Message(content)
        self.acks = 0
        self._queue = queue

    def ack(self):
        self.acks += 1
        self._queue._on_ack(self)


class StubAMQP:
    """in-memory `new AMQP(addr, prefetch, ...)` with listen()/publish() (index.js:43,62,127).

    `prefetch` is AMQP's basic.qos prefetch count: the number of UNACKNOWLEDGED deliveries the broker
    allows; a delivery slot is freed by `ack()`, not by the handler returning.  (In the reference the two
    coincide, because its handler acks before it returns; a handler that defers acks must not rely on it.)"""

    def __init__(self, prefetch: int = 100):
        self.prefetch = prefetch
        self.listeners: Dict[str, Callable] = {}
        self.queues: Dict[str, List[bytes]] = collections.defaultdict(list)
        self.delivered: List[StubRmsg] = []
        self.acked = 0
        self.max_unacked = 0
        self._sem: Optional[asyncio.Semaphore] = None

    def _on_ack(self, rmsg: "StubRmsg"):
        self.acked += 1
        if rmsg.acks == 1 and self._sem is not None:
            self._sem.release()

    async def connect(self):
        return None

    def listen(self, topic: str, fn: Callable):
        self.listeners[topic] = fn

    def publish(self, topic: str, content: bytes):
        self.queues[topic].append(content)

    async def run(self, drain_timeout: Optional[float] = None):
        """deliver everything queued, never more than `prefetch` deliveries unacknowledged; with
        `drain_timeout` also wait (at most that many seconds) until every delivery has been acked, the way a
        consumer keeps running after the queue went quiet"""
        self._sem = sem = asyncio.Semaphore(self.prefetch)
        tasks = []
        for topic, msgs in self.queues.items():
            fn = self.listeners.get(topic)
            if fn is None:
                continue
            for content in msgs:
                await sem.acquire()                     # blocks here while `prefetch` deliveries are unacked
                rmsg = StubRmsg(content, self)
                self.delivered.append(rmsg)
                self.max_unacked = max(self.max_unacked, len(self.delivered) - self.acked)
                tasks.append(asyncio.ensure_future(fn(rmsg)))
            msgs.clear()
        if tasks:
            await asyncio.gather(*tasks)
        if drain_timeout is not None:
            loop = asyncio.get_running_loop()
            t_end = loop.time() + drain_timeout
            while self.acked < len(self.delivered) and loop.time() < t_end:
                await asyncio.sleep(0.001)
        self._sem = None

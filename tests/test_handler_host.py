"""Host logic of the handler mirror (beholder_b200.handler) without a GPU: ack bookkeeping,
deferred acks at commit, side-effect failures — with a fake aggregator standing in for the
C-ABI context (test double only; the product has no such path)."""
import asyncio
import threading

import beholder_b200 as bh
from beholder_b200 import handler as H


class FakeAgg:
    """models the C-ABI context's ack contract: push / submit / poll_committed / flush and the completion
    callback, which fires `latency_s` after the first message of a batch (the library's deadline)"""

    def __init__(self, latency_s=None):
        self.pushed, self.committed, self.submitted = [], 0, 0
        self.latency_s, self.cb, self.timer, self.submits = latency_s, None, None, 0

    def push(self, payload, ts_ms=0):
        self.pushed.append(payload)
        if self.latency_s is not None and self.cb is not None and self.timer is None:
            self.timer = threading.Timer(self.latency_s, self._deadline)
            self.timer.start()

    def _deadline(self):                     # on a foreign thread, like the library's completion thread
        self.timer = None
        self.committed = self.submitted = len(self.pushed)
        self.cb(self.committed)

    def submit(self):
        self.submits += 1
        self.submitted = len(self.pushed)

    def set_commit_callback(self, fn):
        self.cb = fn

    def flush(self):
        self.committed = len(self.pushed)
        return self.committed

    def poll_committed(self):
        self.committed = max(self.committed, self.submitted)      # a submitted batch finishes "at once"
        return self.committed

    def status_counts(self):
        return {"deployed": len(self.pushed)}

    def counters(self):
        return {"dropped_decode": 0, "dropped_enum": 0}

    def close(self):
        pass


def _gpu(commit_every, prefetch=0, latency_s=None):
    g = H.BeholderGPU.__new__(H.BeholderGPU)
    g.schema = bh.telemetry_progress(); g.agg = FakeAgg(latency_s); g.commit_every = commit_every; g.logger = None
    g._init_ack_state(prefetch)
    g.progress_updates_total = H.GpuCounter(g)
    g.trello_comments_total = H.HostCounter("beholder_trello_comments")
    return g


def test_acks_are_deferred_until_commit_and_sent_exactly_once():
    amqp = bh.StubAMQP(prefetch=10)
    g = _gpu(commit_every=4)
    amqp.listen("v1.telemetry.progress", g.handler())
    for i in range(10):
        amqp.publish("v1.telemetry.progress", b"x%d" % i)
    asyncio.run(amqp.run())
    assert amqp.acked == 8                     # two commits of 4; 2 messages still pending
    assert g.commit() == 2
    assert amqp.acked == 10 and all(m.acks == 1 for m in amqp.delivered)
    assert g.commit() == 0


def test_side_effect_errors_are_caught_like_index_js_149():
    amqp = bh.StubAMQP()
    g = _gpu(commit_every=100)
    seen = []

    def side(rmsg):
        seen.append(rmsg.message.content)
        raise ValueError("trello down")
    amqp.listen("v1.telemetry.progress", g.handler(side))
    a = bh.encode(g.schema, mediaId="a", status=4, progress=1); b = bh.encode(g.schema, mediaId="b", status=5, progress=2)
    amqp.publish("v1.telemetry.progress", a); amqp.publish("v1.telemetry.progress", b)
    asyncio.run(amqp.run())
    g.commit()
    assert seen == [a, b] and amqp.acked == 2 and g.agg.pushed == [a, b]


def test_side_effects_only_for_messages_the_reference_would_have_processed():
    # index.js:129-137: a message that does not decode, or whose status has no enum name, lands in the catch
    # before db.getByID / the Trello comment (:140-148): no side effect, still counted as received and acked
    amqp = bh.StubAMQP()
    g = _gpu(commit_every=100)
    seen = []
    amqp.listen("v1.telemetry.progress", g.handler(lambda rmsg: seen.append(rmsg.message.content)))
    good = bh.encode(g.schema, mediaId="m", status=4, progress=7)
    unknown_enum = bh.encode(g.schema, mediaId="m", status=99, progress=7)
    truncated = good[:-1] + b"\x80"                      # the last varint never ends
    bad_wire = b"\x4f" + good                            # unknown field 9 with wire type 7: skipType throws
    for m in (good, unknown_enum, truncated, bad_wire, good):
        amqp.publish("v1.telemetry.progress", m)
    asyncio.run(amqp.run())
    g.commit()
    assert seen == [good, good] and amqp.acked == 5 and len(g.agg.pushed) == 5


def test_host_decoder_agrees_with_the_known_answer_messages():
    import json, os
    from beholder_b200.schema import decode_message
    from beholder_b200 import _lib as L
    kat = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "kat_decode.json")))
    s = bh.telemetry_progress()
    for c in kat["default"]:
        m = decode_message(s, bytes.fromhex(c["payload"]))
        assert (m is None) == (c["outcome"] == L.EV_DROP_DECODE), c["name"]
        if m is not None:
            assert m["status"] == c["status"] and bytes(m["mediaId"]).hex() == c["media_id"], c["name"]
            if c["outcome"] == L.EV_OK:
                assert float(m["progress"]) == c["value"], c["name"]


def test_counter_shape_matches_prom_client():
    g = _gpu(commit_every=100)
    g.agg.pushed = [b"1", b"2"]
    got = g.progress_updates_total.get()
    assert got["name"] == "beholder_progress_updates_total" and got["type"] == "counter"
    assert got["values"] == [{"value": 2, "labels": {"status": "deployed"}}]
    txt = g.metrics()
    assert 'beholder_progress_updates_total{status="deployed"} 2' in txt and "beholder_trello_comments 0" in txt
    try:
        g.progress_updates_total.inc({"status": "x"})
        assert False
    except TypeError:
        pass


def test_prefetch_window_never_stalls_with_default_commit_every():
    # index.js:43: the broker holds back once 100 deliveries are unacked; commit_every (65536) is never reached
    amqp = bh.StubAMQP(prefetch=100)
    g = _gpu(commit_every=65536, prefetch=100)
    amqp.listen("v1.telemetry.progress", g.handler())
    for i in range(10_000):
        amqp.publish("v1.telemetry.progress", b"m%d" % i)
    asyncio.run(asyncio.wait_for(amqp.run(), timeout=20))
    g.commit()
    assert amqp.acked == 10_000 and all(m.acks == 1 for m in amqp.delivered)
    assert amqp.max_unacked <= 100
    assert g.agg.submits >= 10_000 // 50 - 1          # one submit per half prefetch window


def test_a_few_messages_then_idle_are_acked_by_the_completion_callback():
    amqp = bh.StubAMQP(prefetch=100)
    g = _gpu(commit_every=65536, prefetch=100, latency_s=0.005)
    amqp.listen("v1.telemetry.progress", g.handler())
    for i in range(10):
        amqp.publish("v1.telemetry.progress", b"m%d" % i)
    asyncio.run(amqp.run(drain_timeout=5.0))           # nobody calls commit(): the deadline must do it
    assert amqp.acked == 10 and all(m.acks == 1 for m in amqp.delivered)
    assert g.agg.submits == 0


def test_stub_prefetch_counts_unacked_deliveries_not_running_handlers():
    # a handler that returns without acking must block the queue at `prefetch`
    amqp = bh.StubAMQP(prefetch=3)
    got = []

    async def lazy(rmsg):
        got.append(rmsg)
    amqp.listen("q", lazy)
    for i in range(5):
        amqp.publish("q", b"x")

    async def scenario():
        t = asyncio.ensure_future(amqp.run())
        await asyncio.sleep(0.05)
        assert len(got) == 3 and not t.done()          # stuck behind 3 unacked deliveries
        for m in list(got):
            m.ack()
        await asyncio.wait_for(t, timeout=5)
    asyncio.run(scenario())
    assert len(got) == 5

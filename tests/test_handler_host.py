"""Host logic of the handler mirror (beholder_b200.handler) without a GPU: ack bookkeeping,
deferred acks at commit, side-effect failures — with a fake aggregator standing in for the
C-ABI context (test double only; the product has no such path)."""
import asyncio

import beholder_b200 as bh
from beholder_b200 import handler as H


class FakeAgg:
    def __init__(self):
        self.pushed, self.committed = [], 0

    def push(self, payload, ts_ms=0):
        self.pushed.append(payload)

    def flush(self):
        self.committed = len(self.pushed)
        return self.committed

    def poll_committed(self):
        return self.committed

    def status_counts(self):
        return {"deployed": len(self.pushed)}

    def counters(self):
        return {"dropped_decode": 0, "dropped_enum": 0}

    def close(self):
        pass


def _gpu(commit_every):
    g = H.BeholderGPU.__new__(H.BeholderGPU)
    g.schema = bh.telemetry_progress(); g.agg = FakeAgg(); g.commit_every = commit_every; g.logger = None
    import collections
    g._pending = collections.deque(); g._seq = 0
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
    amqp.publish("v1.telemetry.progress", b"a"); amqp.publish("v1.telemetry.progress", b"b")
    asyncio.run(amqp.run())
    g.commit()
    assert seen == [b"a", b"b"] and amqp.acked == 2 and g.agg.pushed == [b"a", b"b"]


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

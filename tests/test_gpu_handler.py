"""The handler mirror on the GPU: BASELINE.json config 1 (reference handler replaying 10k synthetic
events from a local stub queue) — same counts, same /metrics lines, every message acked once."""
import asyncio

import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200.synth import SynthSpec, generate_host
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def test_config1_stub_queue_10k_events():
    spec = SynthSpec(n_keys=64, rate_per_ms=1, malformed_ppm=10000, unknown_enum_ppm=10000)
    slab, off, ts = generate_host(spec, 0, 10_000)
    amqp = bh.StubAMQP(prefetch=100)
    gpu = bh.BeholderGPU(commit_every=2048)
    side = []
    amqp.listen("v1.telemetry.progress", gpu.handler(lambda rmsg: side.append(len(rmsg.message.content))))
    for i in range(10_000):
        amqp.publish("v1.telemetry.progress", bytes(slab[off[i]: off[i + 1]]))
    asyncio.run(amqp.run())
    gpu.commit()
    ref = O.Handler(bh.telemetry_progress())
    ref.handle_slab(slab, off)
    assert gpu.status_counts() == ref.counts()
    assert gpu.dropped() == ref.dropped()
    assert amqp.acked == 10_000 == ref.acks()
    assert all(m.acks == 1 for m in amqp.delivered)
    # the side effects of index.js:140-148 run exactly for the messages the reference counted (decoded, named status)
    assert len(side) == sum(ref.counts().values()) < 10_000
    assert sorted(gpu.progress_updates_total.metrics().splitlines()) == sorted(ref.metrics_text().splitlines())
    vals = {v["labels"]["status"]: v["value"] for v in gpu.progress_updates_total.get()["values"]}
    assert vals == ref.counts()
    gpu.close()


def test_side_effect_failure_does_not_uncount_or_block_ack():
    amqp = bh.StubAMQP()
    gpu = bh.BeholderGPU()

    async def boom(rmsg):
        raise RuntimeError("db down")          # index.js:140 failing lands in the catch at :149
    amqp.listen("v1.telemetry.progress", gpu.handler(boom))
    for st in (4, 4, 5):
        amqp.publish("v1.telemetry.progress", bh.encode(bh.telemetry_progress(), mediaId="m", status=st, progress=3))
    asyncio.run(amqp.run())
    assert gpu.status_counts() == {"deployed": 2, "errored": 1}
    assert amqp.acked == 3
    gpu.close()


def _progress(i):
    return bh.encode(bh.telemetry_progress(), mediaId="m%d" % (i % 7), status=i % 6, progress=i % 101)


def test_prefetch_100_with_default_commit_every_completes():
    """index.js:43 bounds the unacked deliveries at 100: the deferred ack must keep the queue moving although
    commit_every (65536) is never reached"""
    amqp = bh.StubAMQP(prefetch=100)
    gpu = bh.BeholderGPU(prefetch=100)
    amqp.listen("v1.telemetry.progress", gpu.handler())
    msgs = [_progress(i) for i in range(10_000)]
    for m in msgs:
        amqp.publish("v1.telemetry.progress", m)
    asyncio.run(asyncio.wait_for(amqp.run(drain_timeout=10.0), timeout=60))
    assert amqp.acked == 10_000 and all(m.acks == 1 for m in amqp.delivered)
    assert amqp.max_unacked <= 100
    ref = O.Handler(bh.telemetry_progress())
    slab = np.frombuffer(b"".join(msgs), dtype=np.uint8)
    off = np.zeros(len(msgs) + 1, dtype=np.uint32); off[1:] = np.cumsum([len(m) for m in msgs])
    ref.handle_slab(slab, off)
    assert gpu.status_counts() == ref.counts()
    gpu.close()


def test_ten_messages_then_idle_are_acked_within_the_deadline():
    """nobody calls in again after the last message: the library's completion thread submits the partial batch
    at max_batch_latency_ms and reports it; the acks arrive on the event loop"""
    amqp = bh.StubAMQP(prefetch=100)
    gpu = bh.BeholderGPU(prefetch=100, max_batch_latency_ms=5)
    amqp.listen("v1.telemetry.progress", gpu.handler())
    for i in range(10):
        amqp.publish("v1.telemetry.progress", _progress(i))
    import time
    t0 = time.perf_counter()
    asyncio.run(amqp.run(drain_timeout=5.0))
    dt = time.perf_counter() - t0
    assert amqp.acked == 10 and all(m.acks == 1 for m in amqp.delivered)
    assert dt < 1.0, dt                                   # deadline 5 ms + one launch, not the 5 s drain timeout
    assert sum(gpu.agg.status_counts().values()) == 10    # aggregated without any flush from this side
    gpu.close()


def test_poll_committed_submits_a_partial_batch_past_its_deadline():
    import time
    from beholder_b200 import _lib as L
    g = bh.Aggregator(mode=L.MODE_REFERENCE, max_batch_latency_ms=2)
    for i in range(3):
        g.push(_progress(i))
    assert g.poll_committed() == 0                        # younger than the deadline: still staged on the host
    time.sleep(0.01)
    t_end = time.time() + 5
    while g.poll_committed() < 3 and time.time() < t_end:
        time.sleep(0.001)
    assert g.poll_committed() == 3
    never = bh.Aggregator(mode=L.MODE_REFERENCE, max_batch_latency_ms=L.BH_NO_BATCH_DEADLINE)
    never.push(_progress(0)); time.sleep(0.01)
    assert never.poll_committed() == 0
    never.submit()
    t_end = time.time() + 5
    while never.poll_committed() < 1 and time.time() < t_end:
        time.sleep(0.001)
    assert never.poll_committed() == 1
    g.close(); never.close()

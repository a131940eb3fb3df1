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
    assert len(side) == 10_000
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

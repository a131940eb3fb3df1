"""SURVEY §8 f2-f4 on the GPU: /metrics text from the device tables, sibling status stream counting,
snapshot / restore across contexts."""
import asyncio
import os

import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.metrics import render_aggregator, render_windowed
from beholder_b200.synth import SynthSpec, generate_host
from oracle import oracle as O
from helpers import ALL_OPS, SEED, assert_records_equal

pytestmark = pytest.mark.gpu
KEY = ("mediaId", "status")


def test_metrics_text_from_device_tables_equals_oracle_rendering():
    spec = SynthSpec(n_keys=50, rate_per_ms=1)                 # integer progress: sums are exact
    slab, off, ts = generate_host(spec, 0, 30_000)
    ops = ALL_OPS | L.OP_HIST
    with bh.Aggregator(mode=L.MODE_WINDOWED, ops=ops, table_capacity=256) as g:
        g.push_batch(slab, off, ts)
        g.flush()
        got = render_aggregator(g)
    a = O.Agg(bh.telemetry_progress().with_roles(key=KEY), ops, 10_000, SEED)
    a.push_raw(slab, off, ts)
    recs, hist = a.snapshot(with_hist=True)
    assert got == render_windowed(recs.view(bh.RECORD_DTYPE), hist, ops=ops)
    assert "beholder_progress_value_bucket{" in got and 'le="+Inf"' in got


def test_status_stream_counting():
    amqp = bh.StubAMQP()
    gpu = bh.BeholderGPU()
    seen = []
    amqp.listen("v1.telemetry.status", gpu.status_handler(lambda rmsg: seen.append(1)))
    amqp.listen("v1.telemetry.progress", gpu.handler())
    st = bh.telemetry_status()
    for s in (4, 4, 1, 9):                                     # 9 has no enum name
        amqp.publish("v1.telemetry.status", bh.encode(st, mediaId="m", status=s))
    amqp.publish("v1.telemetry.status", b"\x0a\x05ab")         # clamped string, status defaults to 0
    amqp.publish("v1.telemetry.progress", bh.encode(bh.telemetry_progress(), mediaId="m", status=5, progress=1))
    asyncio.run(amqp.run())
    assert gpu.status_counts_sibling() == {"deployed": 2, "downloading": 1, "queued": 1}
    assert gpu.status_counts() == {"errored": 1}
    assert amqp.acked == 6 and len(seen) == 5
    assert 'beholder_status_updates_total{status="deployed"} 2' in gpu.status_updates_total.metrics()
    gpu.close()


@pytest.mark.parametrize("hist", [False, True])
def test_snapshot_restore_across_contexts(tmp_path, hist):
    spec = SynthSpec(n_keys=400, rate_per_ms=2, value_f64=True, value_signed=True, zipf=True)
    slab, off, ts = generate_host(spec, 0, 40_000)
    ops = ALL_OPS | (L.OP_HIST if hist else 0)
    schema = bh.telemetry_progress(L.K_DOUBLE)
    half = 20_000
    path = os.path.join(tmp_path, "state.npz")
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, ops=ops, table_capacity=1024) as g:
        g.push_batch(slab, off[: half + 1], ts[:half])
        g.flush()
        rows = g.save(path)
        assert rows > 400
    with bh.Aggregator(schema, mode=L.MODE_WINDOWED, ops=ops, table_capacity=1024) as g:   # "after the restart"
        assert g.load(path) == rows
        g.push_batch(slab, off[half:], ts[half:])
        g.flush()
        got = g.snapshot_all(with_hist=hist)
    a = O.Agg(schema.with_roles(key=KEY), ops, 10_000, SEED)
    a.push_raw(slab, off, ts)
    if hist:
        want, want_h = a.snapshot(with_hist=True)
        assert_records_equal(got[0], want, ops, got[1], want_h)
    else:
        assert_records_equal(got, a.snapshot(), ops)

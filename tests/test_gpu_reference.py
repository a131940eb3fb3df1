"""GPU parity, reference path (index.js:127-155): per-status counts and drops bit-exact against
the oracle, through the C-ABI (bh_push / bh_push_batch / bh_process_device / bh_snapshot)."""
import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_host, generate_device
from oracle import oracle as O
from helpers import load_json, pack, typed_schema

pytestmark = pytest.mark.gpu
KAT = load_json("kat_decode.json")
EXPECT = load_json("stream_expect.json")


def _gpu_counts(schema, slab, off, **kw):
    with bh.Aggregator(schema, mode=L.MODE_REFERENCE, **kw) as g:
        g.push_batch(slab, off)
        g.flush()
        k = g.counters()
        return g.status_counts(), {"decode": k["dropped_decode"], "enum": k["dropped_enum"]}, k


@pytest.mark.parametrize("group", ["default", "typed"])
def test_kat_vectors_one_by_one(group):
    schema = bh.telemetry_progress() if group == "default" else typed_schema()
    for c in KAT[group]:
        payload = bytes.fromhex(c["payload"])
        with bh.Aggregator(schema, mode=L.MODE_REFERENCE) as g:
            g.push(payload)
            assert g.flush() == 1
            counts = g.status_counts()
            k = g.counters()
        if c["outcome"] == 0:
            assert counts == {schema.enum_name(c["status"]).lower(): 1}, c["name"]
        else:
            assert counts == {}, c["name"]
            assert k["dropped_decode"] == (c["outcome"] == 1) and k["dropped_enum"] == (c["outcome"] == 2), c["name"]


@pytest.mark.parametrize("group", ["default", "typed"])
def test_kat_vectors_as_one_batch(group):
    schema = bh.telemetry_progress() if group == "default" else typed_schema()
    payloads = [bytes.fromhex(c["payload"]) for c in KAT[group]]
    slab, off = pack(payloads)
    counts, dropped, _ = _gpu_counts(schema, slab, off)
    ref = O.Handler(schema)
    ref.handle_slab(slab, off)
    assert counts == ref.counts() and dropped == ref.dropped()


@pytest.mark.parametrize("generic", [False, True])     # the kernel with the reference's schema compiled in / the generic one
@pytest.mark.parametrize("tag", ["c1", "small", "f64_zipf"])
def test_golden_streams(tag, generic):
    e = EXPECT[tag]
    spec = SynthSpec(**e["spec"])
    slab, off, ts = generate_host(spec, 0, e["n"])
    schema = bh.telemetry_progress(L.K_DOUBLE if spec.value_f64 else L.K_INT32)
    counts, dropped, k = _gpu_counts(schema, slab, off, generic_kernels=generic)
    assert counts == e["handler_counts"]
    assert dropped == e["handler_dropped"]
    assert k["events_ok"] + k["dropped_decode"] + k["dropped_enum"] == e["n"]


def test_device_generator_equals_host_generator():
    import torch
    spec = SynthSpec(n_keys=500, rate_per_ms=7, malformed_ppm=20000, unknown_enum_ppm=20000, zipf=True, payload_bytes=80)
    n = 5000
    slab, off, ts = generate_host(spec, 123, n)
    p, o, t, total = generate_device(spec, 123, n)
    torch.cuda.synchronize()
    assert total == int(off[-1])
    assert np.array_equal(o.cpu().numpy().view(np.uint32), off)
    assert np.array_equal(t.cpu().numpy().view(np.uint64), ts)
    assert np.array_equal(p.cpu().numpy()[:total], slab[:total])


@pytest.mark.parametrize("hinted", [False, True])
@pytest.mark.parametrize("payload_bytes", [0, 64, 100, 200, 1024, 4096])
def test_reference_counts_on_device_resident_batches(payload_bytes, hinted):
    import torch
    spec = SynthSpec(n_keys=1000, rate_per_ms=50, malformed_ppm=10000, unknown_enum_ppm=10000, payload_bytes=payload_bytes)
    n = 200_000 if payload_bytes <= 200 else 30_000
    p, o, t, total = generate_device(spec, 0, n)
    # the message-size hint only selects a kernel variant (batches staged whole / in parts): same results
    with bh.Aggregator(mode=L.MODE_REFERENCE, message_bytes_hint=payload_bytes if hinted else 0) as g:
        g.process_device(p, o, None, n)
        g.sync()
        counts = g.status_counts()
        k = g.counters()
    slab = p.cpu().numpy(); off = o.cpu().numpy().view(np.uint32)
    want, dropped = O.run_handler_threads(bh.telemetry_progress(), slab, off, 8)
    assert counts == want
    assert k["dropped_decode"] == int(dropped[1]) and k["dropped_enum"] == int(dropped[2])


def test_push_single_messages_and_slab_rollover():
    spec = SynthSpec(n_keys=30, malformed_ppm=30000, unknown_enum_ppm=30000, payload_bytes=0)
    slab, off, ts = generate_host(spec, 0, 3000)
    ref = O.Handler(bh.telemetry_progress()); ref.handle_slab(slab, off)
    # tiny slabs force many submissions and ring wrap-around
    with bh.Aggregator(mode=L.MODE_REFERENCE, slab_bytes=4096, slab_events=100, n_slabs=2) as g:
        for i in range(3000):
            g.push(bytes(slab[off[i]: off[i + 1]]))
        assert g.flush() == 3000
        assert g.poll_committed() == 3000
        assert g.status_counts() == ref.counts()
    with bh.Aggregator(mode=L.MODE_REFERENCE, slab_bytes=4096, slab_events=64, n_slabs=3) as g:
        g.push_batch(slab, off)
        assert g.flush() == 3000
        assert g.status_counts() == ref.counts()


def test_pinned_batches_and_nonzero_first_offset():
    spec = SynthSpec(n_keys=99, malformed_ppm=10000, payload_bytes=0)
    slab, off, ts = generate_host(spec, 0, 20000)
    ref = O.Handler(bh.telemetry_progress()); ref.handle_slab(slab, off)
    pin_s = bh.PinnedBuffer(len(slab) + 64); pin_o = bh.PinnedBuffer(off.nbytes)
    s = pin_s.view(np.uint8); s[7: 7 + len(slab)] = slab       # unaligned start inside the pinned slab
    o = pin_o.view(np.uint32); o[:] = off + 7
    with bh.Aggregator(mode=L.MODE_REFERENCE, slab_bytes=100_000, slab_events=1500, n_slabs=3) as g:
        g.push_batch(s, o, None, pinned=True)
        assert g.flush() == 20000
        assert g.status_counts() == ref.counts()


def test_empty_and_degenerate_batches():
    with bh.Aggregator(mode=L.MODE_REFERENCE) as g:
        assert g.flush() == 0
        assert g.status_counts() == {}
        slab, off = pack([b"", b"", b""])     # three empty messages: all defaults -> status 0
        g.push_batch(slab, off)
        assert g.flush() == 3
        assert g.status_counts() == {"queued": 3}


def test_oversized_message_uses_global_path():
    big = bh.encode(bh.telemetry_progress(), mediaId="m" * 9000, status=3, progress=5)
    small = bh.encode(bh.telemetry_progress(), mediaId="s", status=4, progress=5)
    slab, off = pack([small, big, small, big[:5000], small])
    ref = O.Handler(bh.telemetry_progress()); ref.handle_slab(slab, off)
    with bh.Aggregator(mode=L.MODE_REFERENCE) as g:
        g.push_batch(slab, off); g.flush()
        # the cut message keeps a clamped mediaId and no status -> proto3 default 0 -> "queued"
        assert g.status_counts() == ref.counts() == {"queued": 1, "uploading": 1, "deployed": 3}


def test_errors_are_reported_not_swallowed():
    with pytest.raises(bh.BeholderError) as e:
        bh.Aggregator(bh.Schema([bh.Field("x", 40, L.K_STRING, 0)], {}), mode=L.MODE_WINDOWED, key=("x",))
    assert e.value.code == L.BH_ERR_UNSUPPORTED
    with bh.Aggregator(mode=L.MODE_REFERENCE, slab_bytes=1024) as g:
        with pytest.raises(bh.BeholderError) as e:
            g.push(b"x" * 2000)
        assert e.value.code == L.BH_ERR_TOO_LARGE

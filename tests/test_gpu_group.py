"""bh_create_group: the key-hash sharded window close inside the library (export rows into the owners' buffers,
flag barrier in peer memory, merge), against the oracle on the whole stream.  The group members may sit on ONE
GPU (plain device pointers instead of peer mappings), which is how a single-GPU box runs these tests; with two or
more GPUs visible the same tests also run across them over NVLink.  The reference has no counterpart: it is one
process with one counter (index.js:23-48,136-138)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.dist import Group, owner_of
from beholder_b200.synth import SynthSpec, generate_device
from oracle import oracle as O
from helpers import ALL_OPS, SEED, assert_records_equal

pytestmark = pytest.mark.gpu
KEY = ("mediaId", "status")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _devices(world):
    import torch
    n = torch.cuda.device_count()
    return [[0] * world] + ([[r % n for r in range(world)]] if n > 1 else [])


def _run(devices, spec, n_per_rank, ops, rounds=2, per_window=False, owner_stride=20_000, cap=1 << 14, flags=0):
    world = len(devices)
    schema = bh.telemetry_progress(L.K_DOUBLE if spec.value_f64 else L.K_INT32)
    hist = bool(ops & L.OP_HIST)
    a = O.Agg(schema.with_roles(key=KEY), ops, 10_000, SEED)
    with Group(devices, owner_stride, schema=schema, mode=L.MODE_WINDOWED, key=KEY, ops=ops, window_ms=10_000,
               table_capacity=cap, window_ring=8, flags=flags) as g:
        first = 0
        for _ in range(rounds):
            for r in range(world):                      # rank r's share of this round: a contiguous run of events
                p, o, t, _ = generate_device(spec, first, n_per_rank, device=devices[r])
                g.local(r).process_device(p, o, t, n_per_rank)
                a.push_raw(p.cpu().numpy(), o.cpu().numpy().view(np.uint32), t.cpu().numpy().view(np.uint64))
                first += n_per_rank
            if per_window:
                wins = sorted(set(w for r in range(world) for w in g.local(r).list_windows()))
                for w in wins:
                    g.close_window(w)
            else:
                g.close_window()
        g.sync()
        parts = [g.owned(r).snapshot_all(with_hist=hist) for r in range(world)]
        drops = [g.local(r).counters() for r in range(world)]
        assert all(len(g.local(r).list_windows()) == 0 for r in range(world))      # every window was recycled
    recs = np.concatenate([p[0] if hist else p for p in parts])
    for r, p in enumerate(parts):
        rows = p[0] if hist else p
        assert (owner_of(rows["key_hash"], world) == r).all()                      # each row sits on its owner
    assert a.hash_collisions() == 0
    if hist:
        want, want_hist = a.snapshot(with_hist=True)
        assert_records_equal(recs, want, ops, np.concatenate([p[1] for p in parts]), want_hist)
    else:
        assert_records_equal(recs, a.snapshot(), ops)
    d = a.dropped()
    assert sum(k["dropped_decode"] for k in drops) == d["decode"] and sum(k["dropped_enum"] for k in drops) == d["enum"]
    assert sum(k["events_ok"] for k in drops) == int(a.snapshot()["count"].sum())


@pytest.mark.parametrize("world", [2, 4])
def test_group_close_equals_oracle(world):
    spec = SynthSpec(n_keys=3000, rate_per_ms=20, malformed_ppm=2000, unknown_enum_ppm=2000)
    for devices in _devices(world):
        _run(devices, spec, 150_000, ALL_OPS)


def test_group_close_with_histograms_window_by_window():
    spec = SynthSpec(n_keys=2000, rate_per_ms=20, value_f64=True, value_exp_lo=-20, value_exp_span=40)
    for devices in _devices(2):
        _run(devices, spec, 120_000, ALL_OPS | L.OP_HIST, rounds=3, per_window=True)


def test_group_on_top_of_the_partitioned_tables():
    spec = SynthSpec(n_keys=60_000, rate_per_ms=100)
    for devices in _devices(2):
        _run(devices, spec, 1_500_000, ALL_OPS, rounds=2, owner_stride=400_000, cap=1 << 17, flags=L.F_PARTITIONED)


def test_cut_rows_are_reported_not_silent():
    spec = SynthSpec(n_keys=3000, rate_per_ms=20)
    with Group([0, 0], 100, mode=L.MODE_WINDOWED, key=KEY, ops=ALL_OPS, table_capacity=1 << 14, window_ring=8) as g:
        for r in range(2):
            p, o, t, _ = generate_device(spec, r * 50_000, 50_000)
            g.local(r).process_device(p, o, t, 50_000)
        g.close_window()
        with pytest.raises(bh.BeholderError) as e:
            g.sync()
        assert e.value.code == L.BH_ERR_TOO_LARGE


@pytest.mark.parametrize("hist", [0, 1])
def test_member_exchange_between_processes_equals_oracle(hist):
    """one process per GPU under torchrun: receive areas travel as CUDA IPC handles, the close is the library's own
    kernels; every rank checks the rows it owns against the oracle on the whole stream"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29549", os.path.join(ROOT, "tools", "probe", "member_exchange_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=dict(os.environ, PEER_HIST=str(hist)))
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    lines = [json.loads(l) for l in out.stdout.splitlines() if l.startswith("{")]
    assert sorted(l["rank"] for l in lines) == [0, 1] and all(l["rows"] > 0 and l["oracle_equal"] for l in lines)

"""world_size-2 (and 4) window-close exchange on CPU over gloo: the host-side logic of
beholder_b200.dist (ownership, fixed-size all-to-all layout) with the oracle standing in for the
per-rank kernels.  Integer aggregates must not depend on the number of ranks."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.dist import all_to_all_records, owner_of, partition_by_owner
from beholder_b200.synth import SynthSpec, generate_host
from oracle import oracle as O
from helpers import ALL_OPS, SEED, assert_records_equal

KEY = ("mediaId", "status")
N_EVENTS = 6000
STRIDE = 1024


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _merge_rows(rows: np.ndarray) -> np.ndarray:
    """owner-side merge of partial rows (what bh_merge_records_device does), via the oracle's merge"""
    schema = bh.telemetry_progress().with_roles(key=KEY)
    total = O.Agg(None, ALL_OPS, 10_000, SEED)
    # feed each partial row as a tiny column aggregator and merge: count/sum add, min/max
    out = {}
    for r in rows:
        if r["count"] == 0:
            continue
        k = (int(r["window_id"]), int(r["key_hash"]))
        if k not in out:
            out[k] = [int(r["count"]), float(r["sum"]), float(r["min"]), float(r["max"])]
        else:
            o = out[k]
            o[0] += int(r["count"]); o[1] += float(r["sum"]); o[2] = min(o[2], float(r["min"])); o[3] = max(o[3], float(r["max"]))
    res = np.zeros(len(out), dtype=O.RECORD_DTYPE)
    for i, ((w, kh), v) in enumerate(sorted(out.items())):
        res[i] = (kh, w, v[0], v[1], v[2], v[3])
    return res


def _worker(rank, world, port, tmp):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    spec = SynthSpec(n_keys=300, rate_per_ms=1, jitter_ms=2000, value_f64=True, value_signed=True)
    slab, off, ts = generate_host(spec, 0, N_EVENTS)
    schema = bh.telemetry_progress(L.K_DOUBLE).with_roles(key=KEY)
    # events by arrival: rank r takes every world-th contiguous block
    lo, hi = N_EVENTS * rank // world, N_EVENTS * (rank + 1) // world
    part = O.Agg(schema, ALL_OPS, 10_000, SEED)
    part.push_raw(slab, off[lo: hi + 1], ts[lo:hi])
    rows = part.snapshot()
    send = partition_by_owner(rows.view(bh.RECORD_DTYPE), world, STRIDE)
    send_t = torch.from_numpy(send.view(np.int64).reshape(world, STRIDE, 6).copy())
    recv_t = all_to_all_records(send_t)
    recv = recv_t.numpy().reshape(-1).view(O.RECORD_DTYPE)
    mine = _merge_rows(recv)
    assert (owner_of(mine["key_hash"], world) == rank).all()
    np.save(os.path.join(tmp, f"rank{rank}.npy"), mine)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_exchange_matches_single_process(world, tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    got = np.concatenate([np.load(tmp_path / f"rank{r}.npy") for r in range(world)])
    spec = SynthSpec(n_keys=300, rate_per_ms=1, jitter_ms=2000, value_f64=True, value_signed=True)
    slab, off, ts = generate_host(spec, 0, N_EVENTS)
    whole = O.Agg(bh.telemetry_progress(L.K_DOUBLE).with_roles(key=KEY), ALL_OPS, 10_000, SEED)
    whole.push_raw(slab, off, ts)
    assert_records_equal(got, whole.snapshot())


def test_owner_and_partition_layout():
    kh = np.array([0, 1 << 63, (1 << 63) - 1, 0xFFFFFFFFFFFFFFFF, 1 << 62], dtype=np.uint64)
    assert owner_of(kh, 1).tolist() == [0, 0, 0, 0, 0]
    assert owner_of(kh, 2).tolist() == [0, 1, 0, 1, 0]
    assert owner_of(kh, 4).tolist() == [0, 2, 1, 3, 1]
    assert owner_of(kh, 8).tolist() == [0, 4, 3, 7, 2]
    recs = np.zeros(5, dtype=bh.RECORD_DTYPE); recs["key_hash"] = kh; recs["count"] = 1
    out = partition_by_owner(recs, 4, 3)
    assert out.shape == (4, 3) and [int((out[o]["count"] > 0).sum()) for o in range(4)] == [1, 2, 1, 1]
    with pytest.raises(ValueError):
        partition_by_owner(recs, 4, 1)

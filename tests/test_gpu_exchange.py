"""Window-close exchange kernels on one GPU: R local contexts stand in for R ranks (each aggregates
its share of the events), their fixed-stride exports are swapped the way the all-to-all would, and
the owners' merged tables must equal the oracle's result over all events — integers bit-exact for
every rank count, sums within 1e-6."""
import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.dist import RECORD_WORDS, owner_of
from beholder_b200.synth import SynthSpec, generate_device
from oracle import oracle as O
from helpers import ALL_OPS, SEED, assert_records_equal

pytestmark = pytest.mark.gpu
KEY = ("mediaId", "status")


@pytest.mark.parametrize("ranks", [1, 2, 4, 8])
@pytest.mark.parametrize("hist", [False, True])
def test_export_swap_merge_equals_oracle(ranks, hist):
    import torch
    n = 240_000
    spec = SynthSpec(n_keys=2000, rate_per_ms=10, value_f64=True, value_signed=True, zipf=True)
    p, o, t, _ = generate_device(spec, 0, n)
    ops = ALL_OPS | (L.OP_HIST if hist else 0)
    schema = bh.telemetry_progress(L.K_DOUBLE)
    mk = lambda: bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ops, table_capacity=4096, window_ring=8)  # noqa: E731
    local = [mk() for _ in range(ranks)]
    owned = [mk() for _ in range(ranks)]
    stride, B = 8192, 64
    off = o.cpu().numpy().view(np.uint32)
    per = n // ranks
    send = []
    for r, g in enumerate(local):
        lo, hi = r * per, (r + 1) * per if r < ranks - 1 else n
        # rank r's share: a contiguous block of events (offsets stay absolute into the shared slab)
        g.process_device(p, o[lo: hi + 1], t[lo:hi], hi - lo)
        recs = torch.zeros(ranks, stride, RECORD_WORDS, dtype=torch.int64, device="cuda")
        hh = torch.zeros(ranks, stride, B, dtype=torch.int32, device="cuda") if hist else None
        counts = torch.zeros(ranks, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        g.export_window_device(L.BH_ALL_WINDOWS, ranks, recs, hh, ranks * stride, counts, owner_stride=stride)
        g.sync()
        assert int(counts.max()) <= stride
        send.append((recs, hh))
    for dst, g in enumerate(owned):       # what all_to_all_single delivers to rank dst: block dst of every sender
        recv = torch.stack([send[src][0][dst] for src in range(ranks)]).contiguous()
        rh = torch.stack([send[src][1][dst] for src in range(ranks)]).contiguous() if hist else None
        torch.cuda.synchronize()
        g.merge_records_device(recv, rh, ranks * stride)
        g.sync()
    got = [g.snapshot_all(with_hist=hist) for g in owned]
    for r, x in enumerate(got):
        rows = x[0] if hist else x
        assert (owner_of(rows["key_hash"], ranks) == r).all()
    recs = np.concatenate([x[0] if hist else x for x in got])
    a = O.Agg.run_threads(schema.with_roles(key=KEY), ops, 10_000, SEED, p.cpu().numpy(), off, t.cpu().numpy().view(np.uint64), 8)
    if hist:
        want, want_h = a.snapshot(with_hist=True)
        assert_records_equal(recs, want, ops, np.concatenate([x[1] for x in got]), want_h)
    else:
        assert_records_equal(recs, a.snapshot(), ops)
    for g in local + owned:
        g.close()

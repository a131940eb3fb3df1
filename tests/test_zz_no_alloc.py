"""bh_export_window_static: the window export for callers that are not allowed to allocate GPU memory
(processes with a frozen memory budget, loops that cannot tolerate cudaMalloc's device-wide
synchronisation).  Its per-owner cursors live in an 8 KB array built into the library, so the call
itself never enters the allocator.

The check: through the C-ABI, the static variant must hand back exactly the rows the regular
bh_export_window_device hands back — same owner counts, same rows per owner, every field bit for bit
(both read the same tables, so even the f64 sums are identical) — on two calls in a row, the second
with a different owner count so that whatever the first call left in the shared array would show.
The file sorts last on purpose: if it fails, a run with -x has already been through every other test.
"""
import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_device

pytestmark = pytest.mark.gpu
ALL_OPS = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX


def _export(agg, fn, n_owners, cap, owner_stride=0, hist=False):
    import torch
    recs = torch.zeros(cap, 6, dtype=torch.int64, device="cuda")
    counts = torch.full((n_owners,), -1, dtype=torch.int64, device="cuda")
    d_hist = torch.zeros(cap, agg.hist_buckets, dtype=torch.int32, device="cuda") if hist else None
    torch.cuda.synchronize()          # the library launches on its own (non-blocking) stream
    fn(L.BH_ALL_WINDOWS, n_owners, recs, d_hist, cap, counts, owner_stride=owner_stride)
    agg.sync()
    torch.cuda.synchronize()
    r = recs.cpu().numpy().reshape(-1).view(bh.RECORD_DTYPE)
    h = d_hist.cpu().numpy().view(np.uint32) if hist else None
    return r, counts.cpu().numpy(), h


def _by_owner(recs, counts, owner_stride, hist=None):
    """rows of each owner, sorted (the order inside an owner's block is unspecified)"""
    out, start = [], 0
    for o, c in enumerate(counts):
        lo = o * owner_stride if owner_stride else start
        rows = recs[lo: lo + c]
        order = np.lexsort((rows["key_hash"], rows["window_id"]))
        out.append((rows[order], hist[lo: lo + c][order] if hist is not None else None))
        start += c
    return out


@pytest.mark.parametrize("hist", [False, True])
def test_static_export_equals_regular_export_twice_in_a_row(hist):
    n = 300_000
    spec = SynthSpec(n_keys=3000, rate_per_ms=10, value_f64=True, value_signed=True)
    p, o, t, _ = generate_device(spec, 0, n)
    ops = ALL_OPS | (L.OP_HIST if hist else 0)
    with bh.Aggregator(bh.telemetry_progress(L.K_DOUBLE), mode=L.MODE_WINDOWED, ops=ops, table_capacity=8192,
                       window_ring=8) as g:
        g.process_device(p, o, t, n)
        g.sync()
        cap = 40_000
        for n_owners, stride in ((8, 0), (2, 0), (4, 9000)):      # consecutive calls share the built-in scratch
            want, want_counts, want_h = _export(g, g.export_window_device, n_owners, cap, stride, hist)
            got, got_counts, got_h = _export(g, g.export_window_static, n_owners, cap, stride, hist)
            assert np.array_equal(got_counts, want_counts)
            assert int(want_counts.sum()) > 3000
            for (a, ah), (b, bh_) in zip(_by_owner(got, got_counts, stride, got_h), _by_owner(want, want_counts, stride, want_h)):
                assert a.tobytes() == b.tobytes()                  # integers, keys, min/max and sums: bit for bit
                if hist:
                    assert np.array_equal(ah, bh_)
            own = got["key_hash"][got["count"] > 0] >> np.uint64(64 - (n_owners.bit_length() - 1))
            assert set(np.unique(own).tolist()) <= set(range(n_owners))


def test_static_export_refuses_more_owners_than_its_scratch_holds():
    import torch
    with bh.Aggregator(mode=L.MODE_WINDOWED, ops=ALL_OPS, table_capacity=64) as g:
        recs = torch.zeros(16, 6, dtype=torch.int64, device="cuda")
        counts = torch.zeros(4096, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        with pytest.raises(bh.BeholderError) as e:
            g.export_window_static(L.BH_ALL_WINDOWS, 2 * L.BH_STATIC_EXPORT_MAX_OWNERS, recs, None, 16, counts)
        assert e.value.code == L.BH_ERR_TOO_LARGE and "built-in scratch" in str(e.value)


def test_static_export_leaves_the_free_device_memory_unchanged():
    """what the entry point exists for, measured: the driver reports the same number of free bytes after the
    calls as before them (outputs allocated beforehand by the caller, as the contract says)"""
    import torch
    n = 100_000
    p, o, t, _ = generate_device(SynthSpec(n_keys=2000, rate_per_ms=10), 0, n)
    with bh.Aggregator(mode=L.MODE_WINDOWED, ops=ALL_OPS, table_capacity=8192, window_ring=8) as g:
        g.process_device(p, o, t, n)
        g.sync()
        recs = torch.zeros(20_000, 6, dtype=torch.int64, device="cuda")
        counts = torch.zeros(8, dtype=torch.int64, device="cuda")
        g.export_window_static(L.BH_ALL_WINDOWS, 8, recs, None, 20_000, counts)     # first call on this device: address look-up
        g.sync(); torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        for n_owners in (8, 4, 1):
            g.export_window_static(L.BH_ALL_WINDOWS, n_owners, recs, None, 20_000, counts)
        g.sync(); torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before
        assert int(counts.cpu()[0]) > 0

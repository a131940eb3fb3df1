"""BASELINE.json config 5 on one GPU: payload 64 B .. 4 KB, HBM-resident (kernel only) vs end to end through
the ingest C-ABI (pinned host -> H2D -> kernel), and config 3's shape (1 M keys, histograms).  One JSON line each."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_device

OPS = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX
st = torch.cuda.Stream(); torch.cuda.set_stream(st); s = st.cuda_stream


def run(payload, n, keys=1000, hist=False, f64=False, cap=2048, e2e=True, zipf=False):
    spec = SynthSpec(n_keys=keys, rate_per_ms=500, payload_bytes=payload, value_f64=f64, zipf=zipf)
    p, o, t, total = generate_device(spec, 0, n)
    ops = OPS | (L.OP_HIST if hist else 0)
    schema = bh.telemetry_progress(L.K_DOUBLE if f64 else L.K_INT32)
    mk = lambda: bh.Aggregator(schema, mode=L.MODE_WINDOWED, ops=ops, table_capacity=cap, window_ring=32 if not hist else 4,  # noqa: E731
                               message_bytes_hint=payload, slab_bytes=64 << 20, slab_events=1 << 20, n_slabs=4)
    g = mk()
    for _ in range(3): g.process_device(p, o, t, n, stream=s)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    e0.record()
    for _ in range(reps): g.process_device(p, o, t, n, stream=s)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    k = g.counters(check=False)
    out = {"payload_bytes": payload, "events": n, "keys": keys, "hist": hist, "f64": f64, "zipf": zipf,
           "bytes_per_event": (total + 12 * n) / n, "hbm_ms": ms, "hbm_events_per_s": n / ms * 1e3,
           "hbm_algorithmic_GBps": (total + 12 * n) / ms / 1e6, "spill_frac": k["smem_spills"] / max(1, k["events_ok"])}
    g.close()
    if e2e:
        pb = bh.PinnedBuffer(p.numel()); ob = bh.PinnedBuffer((n + 1) * 4); tb = bh.PinnedBuffer(n * 8)
        hp, ho, ht = pb.view(np.uint8), ob.view(np.uint32), tb.view(np.uint64)
        torch.from_numpy(hp).copy_(p); torch.from_numpy(ho.view(np.int32)).copy_(o); torch.from_numpy(ht.view(np.int64)).copy_(t)
        torch.cuda.synchronize()
        g = mk()
        g.push_batch(hp, ho, ht, pinned=True); g.flush()
        t0 = time.perf_counter()
        for _ in range(3):
            g.push_batch(hp, ho, ht, pinned=True)
        g.flush()
        dt = (time.perf_counter() - t0) / 3
        out.update({"e2e_events_per_s": n / dt, "e2e_pcie_GBps": (total + 12 * n) / dt / 1e9})
        g.close(); pb.free(); ob.free(); tb.free()
    print(json.dumps(out), flush=True)


for payload in (64, 128, 256, 512, 1024, 2048, 4096):
    run(payload, max(500_000, min(12_500_000, (800 << 20) // payload)))
run(64, 12_500_000, keys=1000, zipf=True, e2e=False)
run(64, 12_500_000, keys=1000, f64=True, e2e=False)
run(64, 12_500_000, keys=1_000_000, hist=True, f64=True, cap=1 << 21, e2e=False)                 # config 3 shape, uniform
run(64, 12_500_000, keys=1_000_000, hist=True, f64=True, cap=1 << 21, e2e=False, zipf=True)      # config 3 shape, Zipf

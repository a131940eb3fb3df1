"""Many-key shapes (BASELINE.json configs 3 and 4) on one GPU, HBM-resident: distinct slabs streamed once, windows
exported to a device buffer and closed as the stream passes them.  One JSON line.
    python tools/c3bench.py [--keys 1000000] [--slabs 8] [--events 12500000] [--no-hist] [--int] [--no-part] [--zipf]
"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_device

ap = argparse.ArgumentParser()
ap.add_argument("--keys", type=int, default=1_000_000); ap.add_argument("--slabs", type=int, default=8)
ap.add_argument("--events", type=int, default=12_500_000, help="per slab")
ap.add_argument("--rate", type=int, default=5000, help="events per ms of stream time")
ap.add_argument("--cap", type=int, default=0); ap.add_argument("--ring", type=int, default=4)
ap.add_argument("--no-hist", action="store_true"); ap.add_argument("--int", action="store_true")
ap.add_argument("--no-part", action="store_true"); ap.add_argument("--zipf", action="store_true")
ap.add_argument("--region", type=int, default=0); ap.add_argument("--replay", type=int, default=0, help="replay the slabs this many more times (same windows)")
ap.add_argument("--no-close", action="store_true")
a = ap.parse_args()

cap = a.cap or (1 << (int(a.keys * 1.5) - 1).bit_length())
spec = SynthSpec(n_keys=a.keys, rate_per_ms=a.rate, jitter_ms=2000, payload_bytes=64, value_f64=not a.int,
                 value_exp_lo=-12, value_exp_span=24, zipf=a.zipf)
ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX | (0 if a.no_hist else L.OP_HIST)
schema = bh.telemetry_progress(L.K_INT32 if a.int else L.K_DOUBLE)
st = torch.cuda.Stream(); torch.cuda.set_stream(st); s = st.cuda_stream
g = bh.Aggregator(schema, mode=L.MODE_WINDOWED, ops=ops, table_capacity=cap, window_ring=a.ring,
                  flags=L.F_NO_PARTITION if a.no_part else 0, part_region_tuples=a.region)
slabs = []
dcdf = None
if a.zipf:
    from beholder_b200.synth import zipf_cdf
    dcdf = torch.from_numpy(zipf_cdf(a.keys).view(np.int64)).cuda()
for i in range(a.slabs):
    p, o, t, total = generate_device(spec, i * a.events, a.events, d_cdf=dcdf)
    slabs.append((p[: (total + 15) // 16 * 16 + 16].clone(), o, t, total))
torch.cuda.synchronize()
out_rows = torch.zeros(cap + 2, 6, dtype=torch.int64, device="cuda")
out_hist = None if a.no_hist else torch.zeros(cap + 2, 64, dtype=torch.int32, device="cuda")
counts = torch.zeros(1, dtype=torch.int64, device="cuda")
WINDOW = 10_000
t0_ms = spec.t0_ms


def run_pass(first_slab_index):
    closed = 0
    open_lo = None
    for i, (p, o, t, total) in enumerate(slabs):
        g.process_device(p, o, t, a.events, stream=s)
        if a.no_close:
            continue
        # watermark: every later event has ts >= t0 + (next index)/rate - jitter
        wm = t0_ms + ((first_slab_index + i + 1) * a.events) // a.rate - 2000
        w_done = wm // WINDOW          # windows < w_done are complete
        if open_lo is None:
            open_lo = (t0_ms - 2000) // WINDOW
        while open_lo < w_done:
            g.export_window_device(open_lo, 1, out_rows, out_hist, cap + 2, counts, stream=s)
            g.close_window_device(open_lo, stream=s)
            open_lo += 1; closed += 1
    return closed


e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
closed = run_pass(0)
g.drain_device(stream=s)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
n = a.slabs * a.events
k = g.counters(check=False)
res = {"args": vars(a), "cap": cap, "events": n, "ms": ms, "events_per_s": n / ms * 1e3, "ms_per_12.5M": ms / n * 12.5e6,
       "algorithmic_GBps": sum(tt + 12 * a.events for *_, tt in slabs) / ms / 1e6, "windows_closed": closed,
       "spill_frac": k["smem_spills"] / max(1, k["events_ok"]), "counters": k}
if a.replay:
    g.reset_device(stream=s); torch.cuda.synchronize()
    run_pass(0); g.drain_device(stream=s); g.reset_device(stream=s); torch.cuda.synchronize()
    e0.record()
    for _ in range(a.replay):
        run_pass(0); g.drain_device(stream=s); g.reset_device(stream=s)
    e1.record(); torch.cuda.synchronize()
    ms2 = e0.elapsed_time(e1) / a.replay
    res.update({"replay_ms": ms2, "replay_events_per_s": n / ms2 * 1e3, "replay_ms_per_12.5M": ms2 / n * 12.5e6})
print(json.dumps(res))

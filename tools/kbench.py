"""Kernel-level timings on one C2 slab (12.5 M events): decode only, columns only, fused.
Usage (GPU box): python tools/kbench.py [--f64] [--zipf] [--keys N] [--events N]"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_device, columns_device

ap = argparse.ArgumentParser()
ap.add_argument("--f64", action="store_true"); ap.add_argument("--zipf", action="store_true")
ap.add_argument("--keys", type=int, default=1000); ap.add_argument("--events", type=int, default=12_500_000)
ap.add_argument("--payload", type=int, default=64); ap.add_argument("--cap", type=int, default=2048)
ap.add_argument("--hist", action="store_true"); ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--unknown-ppm", type=int, default=0); ap.add_argument("--malformed-ppm", type=int, default=0)
a = ap.parse_args()
n = a.events
spec = SynthSpec(n_keys=a.keys, rate_per_ms=500, value_f64=a.f64, zipf=a.zipf, payload_bytes=a.payload,
                 unknown_enum_ppm=a.unknown_ppm, malformed_ppm=a.malformed_ppm)
p, o, t, total = generate_device(spec, 0, n)
kh, ts, val = columns_device(spec, 0xBE401DE5, 0, n)
ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX | (L.OP_HIST if a.hist else 0)
schema = bh.telemetry_progress(L.K_DOUBLE if a.f64 else L.K_INT32)
st = torch.cuda.Stream(); torch.cuda.set_stream(st); s = st.cuda_stream

def timeit(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / a.reps

out = {"events": n, "bytes": total + 12 * n, "args": vars(a)}
g = bh.Aggregator(schema, mode=L.MODE_WINDOWED, ops=ops, table_capacity=a.cap, window_ring=32, message_bytes_hint=a.payload)
dk = torch.empty(n, dtype=torch.int64, device="cuda"); dv = torch.empty(n, dtype=torch.float64, device="cuda")
ds = torch.empty(n, dtype=torch.int32, device="cuda"); do = torch.empty(n, dtype=torch.uint8, device="cuda")
out["decode_ms"] = timeit(lambda: g.decode_device(p, o, n, dk, dv, ds, do, stream=s))
# its own tables: the columnar path has no enum drop, so a stream with unknown statuses would fill them
gc = bh.Aggregator(schema, mode=L.MODE_WINDOWED, ops=ops, table_capacity=a.cap, window_ring=32)
out["columns_ms"] = timeit(lambda: gc.process_columns_device(kh, ts, val, n, stream=s))
out["fused_ms"] = timeit(lambda: g.process_device(p, o, t, n, stream=s))
k = g.counters(check=False); out["spill_frac"] = k["smem_spills"] / max(1, k["events_ok"]); out["counters"] = k; out["slow_frac"] = k["slow_decodes"] / max(1, k["events_ok"])
r = bh.Aggregator(schema, mode=L.MODE_REFERENCE, message_bytes_hint=a.payload)
out["reference_mode_ms"] = timeit(lambda: r.process_device(p, o, None, n, stream=s))
for key in ("decode_ms", "columns_ms", "fused_ms", "reference_mode_ms"):
    out[key.replace("_ms", "_Gev_s")] = n / out[key] / 1e6
out["fused_GBps"] = out["bytes"] / out["fused_ms"] / 1e6
print(json.dumps(out))

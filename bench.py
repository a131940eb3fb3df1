#!/usr/bin/env python
"""bench.py — headline benchmark of the beholder progress-event path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    (N > 1: launched by the driver as `python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...`)

Workload (BASELINE.json configs[1], "C2"): per GPU 100 M synthetic api.TelemetryProgress events of
64 B, 1 000 distinct (mediaId, status) keys, 10 s tumbling windows, count/sum/min/max.  A step is one
pass of the fused decode+aggregate kernel over those events (8 slabs of 12.5 M events = 8 launches),
plus, at N > 1, the window-close exchange of the library (bh_member_close_window: rows stored straight into the key
owners' buffers over NVLink, flag barrier in peer memory, merge; NCCL all-to-all as the fallback).

`value`  : events/s, inputs resident in HBM, CUDA-event timed, max over ranks.
`e2e`    : events/s through the C-ABI ingest (bh_push_batch_compact -> bh_flush -> window close at N > 1 ->
           bh_snapshot): pinned HOST buffers, H2D copies and the D2H read of the result inside the timed region;
           `e2e.variants` (N = 1): the other ingest paths (u32 offsets + u64 timestamps, pageable memory, one call
           per message, buffers registered in place).
`roofline`: fused kernel, algorithmic bytes (payload + 4 B offset + 8 B timestamp per event) over its
           average launch time, against the measured HBM copy bandwidth (MEASURED_PEAKS.json).
`cpu_baseline`: the CPU oracle (a C restatement of the reference handler; Node and triton-core are
           not available, so kind = "port") on the box's host cores, bounded sample.
`c3`, `c5` (N = 1), `c4` (N > 1): BASELINE.json configs 3, 5 and 4 as extra keys of the same line.
`exchange_check` (N > 1): the exchange the timed loop used, against the CPU oracle.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "metric events/sec aggregated"
UNIT = "events/s"
EVENTS_PER_GPU = 100_000_000
SLABS = 8
PAYLOAD_BYTES = 64
N_KEYS = 1000
WINDOW_MS = 10_000
BYTES_PER_EVENT = PAYLOAD_BYTES + 4 + 8      # S + u32 offset + u64 timestamp


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--events", type=int, default=EVENTS_PER_GPU, help="events per GPU per step")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the c3 / c4 / c5 legs (BASELINE.json configs 3-5)")
    ap.add_argument("--c3-events", type=int, default=1_000_000_000, help="events streamed by the c3 leg (per GPU in the c4 leg)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(key="k_process_raw_dram_bytes_per_launch"):
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get(key)
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms from the first warm-up step to the end of the
    timed region (the GPU is under the same load throughout; the timed region alone can be shorter than
    one sampling period)"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx = max(mx, float(r[2]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def spec_for(rank: int, events: int):
    from beholder_b200.synth import SynthSpec
    rate = 500
    # every rank covers the same stream time (round-robin delivery): shift t0 back by its index offset
    first = rank * events
    return SynthSpec(n_keys=N_KEYS, rate_per_ms=rate, jitter_ms=2000, payload_bytes=PAYLOAD_BYTES,
                     t0_ms=1_700_000_000_000 - first // rate), first


def generate_slabs(rank: int, events: int, device: int):
    import torch
    from beholder_b200.synth import generate_device
    spec, first = spec_for(rank, events)
    per = events // SLABS
    slabs = []
    for s in range(SLABS):
        n = per if s < SLABS - 1 else events - per * (SLABS - 1)
        p, o, t, total = generate_device(spec, first + s * per, n, device=device)
        p = p[: (total + 15) // 16 * 16 + 16].clone()     # trim the generator's worst-case capacity
        slabs.append((p, o, t, n, total))
    torch.cuda.synchronize()
    return slabs


# ------------------------------------------------------------------------------------------
# CPU legs (oracle port): cpu_baseline and --impl reference
# ------------------------------------------------------------------------------------------
def host_sample(events: int, device_ok: bool):
    """one slab of the workload in host memory (numpy): from the device generator when a GPU is
    there, else from the host build of the same generator (smaller)."""
    import numpy as np
    if device_ok:
        import torch
        from beholder_b200.synth import generate_device
        spec, first = spec_for(0, EVENTS_PER_GPU)
        p, o, t, total = generate_device(spec, first, events, device=0)
        torch.cuda.synchronize()
        return p[: total + 16].cpu().numpy(), o.cpu().numpy().view(np.uint32), t.cpu().numpy().view(np.uint64)
    from beholder_b200.synth import generate_host
    spec, first = spec_for(0, EVENTS_PER_GPU)
    return generate_host(spec, first, events)


def cpu_windowed(slab, off, ts, threads: int):
    import beholder_b200 as bh
    from beholder_b200 import _lib as L
    from oracle import oracle as O
    schema = bh.telemetry_progress().with_roles(key=("mediaId", "status"))
    ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX
    t0 = time.perf_counter()
    a = O.Agg.run_threads(schema, ops, WINDOW_MS, 0xBE401DE5, slab, off, ts, threads)
    dt = time.perf_counter() - t0
    rows = len(a.snapshot())
    return (len(off) - 1) / dt, rows


def cpu_baseline_block(device_ok: bool):
    import beholder_b200 as bh
    from oracle import oracle as O
    cores = os.cpu_count() or 1
    n = 12_500_000 if device_ok else 200_000
    slab, off, ts = host_sample(n, device_ok)
    v_all, rows = cpu_windowed(slab, off, ts, cores)
    m = min(n, 2_000_000)
    v_one, _ = cpu_windowed(slab, off[: m + 1], ts[:m], 1)
    out = {"value": v_all, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": f"{n} events = 1 of {SLABS} slabs of the C2 workload (windowed count/sum/min/max, oracle C port, "
                     f"{cores} threads with private tables + merge); the reference itself (Node + triton-core) cannot run here",
           "single_thread_value": v_one, "single_thread_sample_events": m, "rows": rows}
    # the reference handler proper (count by status, index.js:127-155), 1 thread like Node, logger on/off
    h = O.Handler(bh.telemetry_progress(), 0)
    k = min(n, 2_000_000)
    t0 = time.perf_counter(); O.run_handler_threads(bh.telemetry_progress(), slab, off[: k + 1], 1, 0); d0 = time.perf_counter() - t0
    t0 = time.perf_counter(); O.run_handler_threads(bh.telemetry_progress(), slab, off[: k + 1], 1, 1); d1 = time.perf_counter() - t0
    out["reference_handler_1thread_log_off"] = k / d0
    out["reference_handler_1thread_log_on"] = k / d1
    del h
    return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (the oracle port: Node and
    triton-core are unavailable) on all host threads, same metric/config, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    try:
        import torch
        device_ok = torch.cuda.is_available()
    except Exception:
        device_ok = False
    cores = os.cpu_count() or 1
    n = 12_500_000 if device_ok else 200_000
    slab, off, ts = host_sample(n, device_ok)
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        cpu_windowed(slab, off, ts, cores)
        if i >= args.warmup:
            times.append(time.perf_counter() - t0)
    dt = sum(times) / len(times)
    v = n / dt
    sample = (f"each step = {n} events (1 of {SLABS} slabs of the C2 workload) through the oracle C port of "
              f"index.js:127-155 + §B.2 on {cores} threads")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64 counts / f64 sum,min,max", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def oracle_exchange_check(local, owned, exch, world, rank, device, stream, n=1_000_000):
    """The exchange the timed loop used, against the CPU oracle: every rank feeds the first n events of its share,
    one window close moves the rows to their owners, and each rank compares the rows it now owns with the oracle's
    table of ALL ranks' events reduced to its keys.  Outside every timed region."""
    import numpy as np
    import torch
    import beholder_b200 as bh
    from beholder_b200 import _lib as L
    from beholder_b200.dist import owner_of
    from beholder_b200.synth import generate_device
    from oracle import oracle as O
    torch.cuda.synchronize()
    local.reset_device(stream=stream); owned.reset_device(stream=stream)
    spec, first = spec_for(rank, EVENTS_PER_GPU)
    p, o, t, _ = generate_device(spec, first, n, device=device)
    local.process_device(p, o, t, n, stream=stream)
    exch.close_all_windows(reset_local=True)
    torch.cuda.synchronize()
    owned.sync()
    got = owned.snapshot_all()
    schema = bh.telemetry_progress().with_roles(key=("mediaId", "status"))
    ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX
    a = O.Agg(schema, ops, WINDOW_MS, 0xBE401DE5)
    for r in range(world):
        sp, fr = spec_for(r, EVENTS_PER_GPU)
        p, o, t, _ = generate_device(sp, fr, n, device=device)
        part = O.Agg.run_threads(schema, ops, WINDOW_MS, 0xBE401DE5, p.cpu().numpy(), o.cpu().numpy().view(np.uint32),
                                 t.cpu().numpy().view(np.uint64), min(8, os.cpu_count() or 1))
        a.merge(part)
    want = a.snapshot()
    want = want[owner_of(want["key_hash"], world) == rank]
    got = got[np.lexsort((got["key_hash"], got["window_id"]))]
    want = want[np.lexsort((want["key_hash"], want["window_id"]))]
    assert len(got) == len(want) and len(want) > 0, (len(got), len(want))
    for f in ("key_hash", "window_id", "count"):
        assert np.array_equal(got[f], want[f]), f
    for f in ("min", "max"):
        assert np.array_equal(got[f].view(np.uint64), want[f].view(np.uint64)), f
    assert np.allclose(got["sum"], want["sum"], rtol=1e-6, atol=0)
    local.reset_device(stream=stream); owned.reset_device(stream=stream)
    torch.cuda.synchronize()
    return "oracle-equal"


def workload_config(n_gpus: int):
    return {"workload": f"C2: per GPU {EVENTS_PER_GPU} synthetic api.TelemetryProgress events x {PAYLOAD_BYTES} B, "
                        f"{N_KEYS} distinct (mediaId,status) keys, {WINDOW_MS // 1000} s tumbling windows, count/sum/min/max, "
                        f"raw protobuf payloads decoded on the GPU",
            "events_per_gpu": EVENTS_PER_GPU, "slabs_per_step": SLABS, "payload_bytes": PAYLOAD_BYTES, "keys": N_KEYS,
            "window_ms": WINDOW_MS, "ops": "count,sum,min,max", "key_distribution": "uniform",
            "l2_policy": "inputs (7.6 GB per GPU per step) are larger than L2; no explicit flush",
            "parallelism": f"events sharded over {n_gpus} GPU(s); key-hash owner merge per step when N>1 "
                           f"(export kernel stores into the owners' buffers over NVLink, or NCCL all-to-all: see 'exchange')"}


# ------------------------------------------------------------------------------------------
# BASELINE.json configs 3, 4, 5 beside the headline: one dictionary each in the JSON line
# ------------------------------------------------------------------------------------------
MANY_RATE = 5000            # events per ms of stream time: 1 B events = 200 s = 20 windows of 10 s
MANY_SLAB = 12_500_000
MANY_JITTER = 2000


def many_key_pass(local, slabs, stream, on_close):
    """one pass over pre-generated slabs of a many-key stream, windows closed behind the watermark (every rank's
    slabs cover the same stream time: many_key_slabs shifts t0 by the rank's first event index)"""
    import beholder_b200 as bh
    t0 = 1_700_000_000_000
    ws = bh.WindowStream(local, WINDOW_MS, on_close)
    g = 0
    for p, o, t, n, _ in slabs:
        ws.process_device(p, o, t, n, first_ts_ms=t0 + g // MANY_RATE - MANY_JITTER,
                          watermark_ms=t0 + (g + n) // MANY_RATE - MANY_JITTER, stream=stream)
        g += n
    ws.finish(t0 + g // MANY_RATE + MANY_JITTER, stream)
    return ws.closed


def many_key_slabs(keys, events, rank, device, f64, zipf=False):
    import numpy as np
    import torch
    from beholder_b200.synth import SynthSpec, generate_device, zipf_cdf
    spec = SynthSpec(n_keys=keys, rate_per_ms=MANY_RATE, jitter_ms=MANY_JITTER, payload_bytes=PAYLOAD_BYTES, value_f64=f64,
                     value_exp_lo=-12, value_exp_span=24, zipf=zipf)
    d_cdf = torch.from_numpy(zipf_cdf(keys).view(np.int64)).to(torch.device("cuda", device)) if zipf else None
    slabs, g = [], 0
    first = rank * events           # ranks take consecutive stretches of the event index space ...
    spec.t0_ms -= first // MANY_RATE  # ... that cover the same stream time (round-robin delivery)
    while g < events:
        n = min(MANY_SLAB, events - g)
        p, o, t, total = generate_device(spec, first + g, n, device=device, d_cdf=d_cdf)
        slabs.append((p[: (total + 15) // 16 * 16 + 16].clone(), o, t, n, total))
        g += n
    torch.cuda.synchronize()
    return slabs, first


def _c3_pass(events, f64, device, stream, zipf=False):
    import torch
    import beholder_b200 as bh
    from beholder_b200 import _lib as L
    keys, cap = 1_000_000, 1 << 21
    ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX | L.OP_HIST
    slabs, first = many_key_slabs(keys, events, 0, device, f64, zipf)
    g = bh.Aggregator(bh.telemetry_progress(L.K_DOUBLE if f64 else L.K_INT32), mode=L.MODE_WINDOWED, ops=ops, window_ms=WINDOW_MS,
                      table_capacity=cap, window_ring=4, device=device)
    rows = torch.zeros(cap + 2, 6, dtype=torch.int64, device="cuda")
    hist = torch.zeros(cap + 2, 64, dtype=torch.int32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    exported = torch.zeros(1, dtype=torch.int64, device="cuda")

    def on_close(w, st):
        g.export_window_device(w, 1, rows, hist, cap + 2, cnt, stream=st)
        exported.add_(cnt)
        g.close_window_device(w, stream=st)

    # warm-up (timing rule: never time a kernel's first launches): the first slab three times, windows closed
    for _ in range(3):
        many_key_pass(g, slabs[:1], stream, on_close)
    g.reset_device(stream=stream); exported.zero_()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    closed = many_key_pass(g, slabs, stream, on_close)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    k = g.counters()
    assert k["events_ok"] == events and k["dropped_table_full"] == 0 and k["dropped_window"] == 0, k
    assert not g.list_windows()
    nbytes = sum(total + 12 * n for *_, n, total in slabs)
    g.close()
    del slabs, rows, hist
    torch.cuda.empty_cache()
    return {"value": events / (ms * 1e-3), "ms": ms, "events": events, "windows_closed": closed, "rows_exported": int(exported.item()),
            "smem_miss_frac": k["smem_spills"] / events, "gbs": nbytes / (ms * 1e-3) / 1e9}


def c3_leg(args, device, stream, peak):
    """configs[2]: 1 GPU, 1 B events streamed once, 1 M distinct keys, 64-bucket log2 histograms, f64 values.  The
    whole stream is resident in HBM (76 GB of the 180); tables live in a ring of 4 windows, every window is
    exported to a device buffer (rows + histograms) and recycled as the watermark passes it."""
    r = _c3_pass(args.c3_events, True, device, stream)
    out = {"workload": f"C3: {r['events']} events x {PAYLOAD_BYTES} B streamed once, 1000000 distinct keys, {WINDOW_MS // 1000} s windows, "
                       f"count/sum/min/max + 64-bucket log2 histograms, f64 values; window ring 4, each window exported (rows + "
                       f"histograms) to a device buffer and recycled behind the watermark",
           "value": r["value"], "unit": UNIT, "ms": r["ms"], "events": r["events"], "keys": 1_000_000, "windows_closed": r["windows_closed"],
           "rows_exported": r["rows_exported"], "smem_miss_frac": r["smem_miss_frac"],
           "roofline": {"bound": "hbm", "kernels": "bh::k_process_raw + bh::k_part_drain + export/close", "achieved": r["gbs"],
                        "peak": peak, "unit": "GB/s", "frac": r["gbs"] / peak, "algorithmic_bytes_per_event": BYTES_PER_EVENT,
                        "traffic_bytes_per_event": ncu_traffic("c3_dram_bytes_per_event"),
                        "note": "whole pass (decode + log + fold + export + close) against the 76 B/event the input costs; table "
                                "and log traffic are extra (traffic_bytes_per_event, from ncu)"}}
    # the same shape with the reference's own int32 `progress` (the schema-specialised fused kernel, integer fold)
    v = _c3_pass(min(args.c3_events, 200_000_000), False, device, stream)
    out["variant_int32_progress"] = {"value": v["value"], "unit": UNIT, "ms": v["ms"], "events": v["events"], "frac": v["gbs"] / peak,
                                     "windows_closed": v["windows_closed"]}
    # Zipf(1.0) keys (SURVEY 8d asks for both distributions): hot keys stay in the per-SM table, the tail is logged
    z = _c3_pass(min(args.c3_events, 200_000_000), True, device, stream, zipf=True)
    out["variant_zipf_keys"] = {"value": z["value"], "unit": UNIT, "ms": z["ms"], "events": z["events"], "frac": z["gbs"] / peak,
                                "smem_miss_frac": z["smem_miss_frac"], "windows_closed": z["windows_closed"]}
    return out


def c4_leg(args, world, rank, device, stream, peak):
    """configs[3]: N GPUs, events spread by arrival, 10 M distinct keys seen by every GPU, per-window partial
    aggregates moved to their key-hash owners over NVLink and merged there (the library's own exchange)."""
    import torch
    import torch.distributed as dist
    import beholder_b200 as bh
    from beholder_b200 import _lib as L
    from beholder_b200.dist import MemberExchange
    events, keys = args.c3_events, 10_000_000
    ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX
    slabs, first = many_key_slabs(keys, events, rank, device, False)
    local = bh.Aggregator(mode=L.MODE_WINDOWED, ops=ops, window_ms=WINDOW_MS, table_capacity=1 << 24, window_ring=4, device=device)
    own_cap = 1 << max(16, (int(keys * 1.6 / world) - 1).bit_length())
    owned = bh.Aggregator(mode=L.MODE_WINDOWED, ops=ops, window_ms=WINDOW_MS, table_capacity=own_cap, window_ring=32, device=device,
                          flags=L.F_NO_PARTITION)
    stride = int(keys / world * 1.15) + 4096
    ex = MemberExchange(local, owned, world, rank, owner_stride=stride, device=device)
    t_close = []

    def on_close(w, st):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); ex.close_window(w, st); b.record()
        t_close.append((a, b))

    for _ in range(3):                                # warm-up: first slab, windows closed through the exchange
        many_key_pass(local, slabs[:1], stream, on_close)
    torch.cuda.synchronize(); dist.barrier()
    local.reset_device(stream=stream); owned.reset_device(stream=stream); t_close.clear()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    closed = many_key_pass(local, slabs, stream, on_close)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    sent = ex.check_no_overflow()                     # rows of the last close per owner; raises if any block was cut
    k = local.counters()
    assert k["events_ok"] == events and k["dropped_table_full"] == 0 and k["dropped_window"] == 0, k
    owned.sync()
    ko = owned.counters()
    assert ko["dropped_table_full"] == 0 and ko["dropped_window"] == 0, ko
    # every event counted exactly once across the owners' tables
    total_owned = 0
    for w in owned.list_windows():
        total_owned += int(owned.snapshot(w, cap=own_cap + 2)["count"].sum())
    close_ms = sum(a.elapsed_time(b) for a, b in t_close)
    t = torch.tensor([ms, close_ms, float(total_owned), float(sum(sent))], dtype=torch.float64, device="cuda")
    tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    ms, close_ms = tmax[0].item(), tmax[1].item()
    assert int(tsum[2].item()) == world * events, (int(tsum[2].item()), world * events)
    rows_per_close = tsum[3].item() / world
    nvlink_bytes = rows_per_close * 48 * (world - 1) / world
    out = {"workload": f"C4: {world} GPUs x {events} events x {PAYLOAD_BYTES} B, {keys} distinct keys seen by every GPU, {WINDOW_MS // 1000} s "
                       f"windows, count/sum/min/max; per-window partial rows stored into their key-hash owners' buffers over NVLink and "
                       f"merged (bh_member_close_window), window ring 4",
           "value": world * events / (ms * 1e-3), "unit": UNIT, "ms": ms, "events_per_gpu": events, "keys": keys,
           "windows_closed": closed, "close_ms_total": close_ms, "close_ms_per_window": close_ms / max(1, closed),
           "close_share_of_pass": close_ms / ms, "rows_sent_per_gpu_per_close": rows_per_close,
           "nvlink_bytes_per_gpu_per_close": nvlink_bytes,
           "nvlink_gbs_during_close": nvlink_bytes / (close_ms / max(1, closed) * 1e-3) / 1e9 if close_ms else None,
           "exchange_check": "every event counted once in the owners' tables (sum of counts == N x events)",
           "per_gpu_algorithmic_gbs": events * BYTES_PER_EVENT / (ms * 1e-3) / 1e9, "frac_of_hbm_peak": events * BYTES_PER_EVENT / (ms * 1e-3) / 1e9 / peak}
    ex.close(); local.close(); owned.close()
    del slabs
    torch.cuda.empty_cache()
    return out


def c5_leg(device, stream, peak, pcie_peak, world=1, rank=0):
    """configs[4]: payload 64 B .. 4 KB on every GPU of the run at once, HBM-resident and end to end; events/s are
    summed over the GPUs, fractions are per GPU (HBM) and against `pcie_peak` (the link peak the caller measured:
    one GPU alone at N = 1, all GPUs copying at once at N > 1).  Bytes: `hbm_read_gbs` counts what the kernel has to
    touch -- all of a message while batches are staged through shared memory (<= 288 B), the fields it parses plus
    the headers it skips once messages are read from global memory (the pad is skipped)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import beholder_b200 as bh

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def summed(v, op="sum"):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX)
        return t.item()

    from beholder_b200 import _lib as L
    from beholder_b200.synth import SynthSpec, generate_device
    ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX
    out = []
    for payload in (64, 128, 256, 512, 1024, 2048, 4096):
        n = max(500_000, min(12_500_000, (800 << 20) // payload))
        spec = SynthSpec(n_keys=N_KEYS, rate_per_ms=500, payload_bytes=payload)
        p, o, t, total = generate_device(spec, rank * n, n, device=device)
        mk = lambda: bh.Aggregator(mode=L.MODE_WINDOWED, ops=ops, window_ms=WINDOW_MS, table_capacity=2048, window_ring=32,  # noqa: E731
                                   device=device, message_bytes_hint=payload, slab_bytes=64 << 20, slab_events=1 << 20, n_slabs=4)
        g = mk()
        for _ in range(3):
            g.process_device(p, o, t, n, stream=stream)
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            g.process_device(p, o, t, n, stream=stream)
        e1.record(); torch.cuda.synchronize()
        ms = summed(e0.elapsed_time(e1) / 5, "max")
        g.close()
        staged = payload <= 288
        touched = (total if staged else n * 96) + 12 * n     # global-memory reader: ~64 B of fields + 32 B sectors of the skipped pad's header
        pb = bh.PinnedBuffer(p.numel()); ob = bh.PinnedBuffer((n + 1) * 4); tb = bh.PinnedBuffer(n * 8)
        hp, ho, ht = pb.view(np.uint8), ob.view(np.uint32), tb.view(np.uint64)
        torch.from_numpy(hp).copy_(p); torch.from_numpy(ho.view(np.int32)).copy_(o); torch.from_numpy(ht.view(np.int64)).copy_(t)
        torch.cuda.synchronize()
        g = mk()
        g.push_batch(hp, ho, ht, pinned=True); g.flush()
        sync_all()
        t0 = time.perf_counter()
        for _ in range(2):
            g.push_batch(hp, ho, ht, pinned=True)
        g.flush()
        sync_all()
        dt = summed((time.perf_counter() - t0) / 2, "max")
        g.close(); pb.free(); ob.free(); tb.free()
        wire = total + 12 * n
        hbm_frac, pcie_frac = touched / (ms * 1e-3) / 1e9 / peak, world * wire / dt / 1e9 / pcie_peak
        out.append({"payload_bytes": payload, "events_per_gpu": n, "hbm_events_per_s": world * n / (ms * 1e-3), "hbm_read_gbs_per_gpu": touched / (ms * 1e-3) / 1e9,
                    "hbm_frac": hbm_frac, "e2e_events_per_s": world * n / dt, "e2e_pcie_gbs": world * wire / dt / 1e9, "e2e_pcie_frac": pcie_frac,
                    "regime": "ingest-bound end to end (PCIe); HBM-resident: " +
                              ("instruction issue (decode)" if staged else "latency of per-message global reads (pad skipped, not read)")})
        del p, o, t
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------
def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    import beholder_b200 as bh
    from beholder_b200 import _lib as L
    from beholder_b200.dist import WindowExchange

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched through torch.distributed.run (one rank per GPU)")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: beholder_b200 has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        # keep stdout to the one JSON line: NCCL prints its version banner there at NCCL_DEBUG=VERSION
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    events = args.events
    ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX
    mk = lambda: bh.Aggregator(mode=L.MODE_WINDOWED, ops=ops, window_ms=WINDOW_MS, table_capacity=2048,  # noqa: E731
                               window_ring=32, device=local_rank, slab_bytes=64 << 20, slab_events=1 << 20, n_slabs=4)
    local = mk()
    slabs = generate_slabs(rank, events, local_rank)
    # everything timed runs on ONE explicit torch stream: our launches (handle passed through the
    # C-ABI), NCCL, and the CUDA events that time them
    tstream = torch.cuda.Stream(device=local_rank)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0
    exch = None
    if world > 1:
        owned = mk()
        stride = 2 * (N_KEYS * 40) // world
        # "member" (default): the library's own exchange -- receive areas shared as CUDA IPC handles, the export
        # kernel stores rows straight into the owners' buffers over NVLink, a flag barrier in peer memory, merge;
        # torch.distributed only carried the handles.  "nccl": export -> all_to_all_single -> merge.  A failure to
        # map the peers' memory raises on every rank together (MemberExchange), so all of them fall back together.
        exchange = os.environ.get("BEHOLDER_EXCHANGE", "member")
        if exchange == "member":
            try:
                from beholder_b200.dist import MemberExchange
                exch = MemberExchange(local, owned, world, rank, owner_stride=stride, device=local_rank)
            except Exception as e:      # noqa: BLE001 - CUDA IPC between the ranks unavailable on this box
                print(f"[bench] rank {rank}: member exchange unavailable ({e}); using NCCL all-to-all", file=sys.stderr)
                exch = None
        if exch is None:
            exchange = "nccl"
            exch = WindowExchange(local, owned, world, rank, owner_stride=stride, device=local_rank)

    def step(ev_mid=None):
        for p, o, t, n, _ in slabs:
            local.process_device(p, o, t, n, stream=stream)
        if ev_mid is not None:
            ev_mid.record()
        if exch is not None:
            exch.close_all_windows(reset_local=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)     # let nvidia-smi come up before the first warm-up step
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    barrier()
    e_start, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e_start.record()
    for a, m, b in ev:
        a.record(); step(m); b.record()
    e_end.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    total_ms = e_start.elapsed_time(e_end)
    kernel_ms = sum(a.elapsed_time(m) for a, m, _ in ev)
    t = torch.tensor([total_ms, kernel_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, kernel_ms = t.tolist()
    value = world * events * args.steps / (total_ms * 1e-3)

    # correctness guard on what was just timed: every event counted exactly once
    exchange_check = None
    if exch is None:
        k = local.counters()
        done = max(args.warmup, 3) + args.steps
        assert k["events_ok"] == events * done, (k, events * done)
    else:
        exch.check_no_overflow()
        k = local.counters()
        assert k["dropped_table_full"] == 0 and k["dropped_window"] == 0, k
        done = max(args.warmup, 3) + args.steps
        if exchange == "member":       # (the NCCL exchange resets the local counters with the tables)
            assert k["events_ok"] == events * done, (k, events * done)
        # every event of every rank and step sits in exactly one owner's table
        owned.sync()
        mine = sum(int(owned.snapshot(w, cap=4096)["count"].sum()) for w in owned.list_windows())
        tot = torch.tensor([float(mine)], dtype=torch.float64, device="cuda")
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        assert int(tot.item()) == world * events * done, (int(tot.item()), world * events * done)
        exchange_check = oracle_exchange_check(local, owned, exch, world, rank, local_rank, stream)

    # ---- the reference's literal semantics on the same slabs (index.js:136-138: one counter per status,
    # no key hash, no windows, no timestamps): reported beside the headline, not instead of it ----
    ref_mode = None
    if world == 1 and not args.no_e2e:
        rg = bh.Aggregator(mode=L.MODE_REFERENCE, device=local_rank)
        for _ in range(3):
            for p, o, t_, n, _ in slabs:
                rg.process_device(p, o, None, n, stream=stream)
        torch.cuda.synchronize()
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        rsteps = min(args.steps, 10)
        r0.record()
        for _ in range(rsteps):
            for p, o, t_, n, _ in slabs:
                rg.process_device(p, o, None, n, stream=stream)
        r1.record(); torch.cuda.synchronize()
        rms = r0.elapsed_time(r1) / rsteps
        rk = rg.counters()
        assert rk["events_ok"] == events * (3 + rsteps), rk
        rbytes = sum(total + (n + 1) * 4 for *_, n, total in slabs)      # payload + offsets; no timestamps in this mode
        ref_mode = {"value": events / (rms * 1e-3), "unit": UNIT, "ms_per_step": rms, "steps": rsteps,
                    "what": "BH_MODE_REFERENCE: beholder_progress_updates_total{status} counts only (index.js:30-34,136-138)",
                    "algorithmic_bytes_per_event": rbytes / events, "achieved_gbs": rbytes / (rms * 1e-3) / 1e9,
                    "status_counts": rg.status_counts()}
        rg.close()

    # ---- end to end through the ingest C-ABI: pinned host buffers -> H2D -> kernel -> D2H result ----
    e2e = None
    if not args.no_e2e:
        pins = []
        for p, o, t_, n, total in slabs:
            pb = bh.PinnedBuffer(p.numel()); ob = bh.PinnedBuffer((n + 1) * 4); tb = bh.PinnedBuffer(n * 8)
            hp, ho, ht = pb.view(np.uint8), ob.view(np.uint32), tb.view(np.uint64)
            torch.from_numpy(hp).copy_(p); torch.from_numpy(ho.view(np.int32)).copy_(o); torch.from_numpy(ht.view(np.int64)).copy_(t_)
            pins.append((pb, ob, tb, hp, ho, ht, n, total))
        # compact sidecars of the same slabs (u16 lengths, i16 timestamp deltas, one offset + base per 65 536 events):
        # what the end-to-end leg sends, 68 instead of 76 bytes per event over the link that bounds it
        from beholder_b200.engine import compact_sidecars
        cpins = []
        for _, _, _, hp, ho, ht, n, total in pins:
            arrs = compact_sidecars(ho, ht)
            bufs = [bh.PinnedBuffer(a.nbytes + 64) for a in arrs]
            views = []
            for b, a in zip(bufs, arrs):
                v = b.view(a.dtype, count=len(a)); v[:] = a; views.append(v)
            cpins.append((bufs, views, hp[: total + 16], n, total))
        torch.cuda.synchronize()
        # the ingest roofline: plain pinned-host -> device copies of the first slab (CUDA events).
        # One GPU alone (best of 5); with N > 1 also all N GPUs copying at the same time (5 copies back to
        # back, summed over ranks) -- on a box whose GPUs share PCIe uplinks / one host memory system the
        # concurrent figure, not N x the solo one, is the ceiling of the N-GPU ingest.
        src = torch.from_numpy(pins[0][3]); dst = torch.empty_like(slabs[0][0])

        def copy_gbs(reps, best_of):
            rates = []
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if not best_of:
                c0.record()
            for _ in range(reps):
                if best_of:
                    c0.record()
                dst.copy_(src, non_blocking=True)
                if best_of:
                    c1.record(); torch.cuda.synchronize()
                    rates.append(src.numel() / (c0.elapsed_time(c1) * 1e-3) / 1e9)
            if best_of:
                return max(rates)
            c1.record(); torch.cuda.synchronize()
            return reps * src.numel() / (c0.elapsed_time(c1) * 1e-3) / 1e9

        pcie_concurrent = None
        if world > 1:
            barrier()
            solo = copy_gbs(5, True) if rank == 0 else 0.0      # the other ranks wait at the barrier
            barrier()
            mine = copy_gbs(5, False)
            conc = torch.tensor([mine, solo], dtype=torch.float64, device="cuda")
            dist.all_reduce(conc, op=dist.ReduceOp.SUM)
            pcie_concurrent_sum, pcie_peak = conc[0].item(), conc[1].item()
            # same convention as every other number here: the job is done when the LAST rank is done
            slowest = torch.tensor([5 * src.numel() / (mine * 1e9)], dtype=torch.float64, device="cuda")   # seconds
            dist.all_reduce(slowest, op=dist.ReduceOp.MAX)
            pcie_concurrent = world * 5 * src.numel() / slowest.item() / 1e9
            barrier()
        else:
            pcie_peak = copy_gbs(5, True)
        del dst
        ing = mk()
        h2d_plain = sum(total + (n + 1) * 4 + n * 8 for *_, n, total in pins)
        h2d = sum(total + sum(v.nbytes for v in views) for _, views, _, n, total in cpins)
        ing_owned = ing_ex = None
        if world > 1 and exchange == "member":
            from beholder_b200.dist import MemberExchange
            ing_owned = mk()
            ing_ex = MemberExchange(ing, ing_owned, world, rank, owner_stride=stride, device=local_rank)

        def e2e_step(compact=True):
            if compact:
                for _, (hl, hb, htb, htd), hpc, n, _ in cpins:
                    ing.push_batch_compact(hpc, hl, hb, htb, htd)
            else:
                for _, _, _, hp, ho, ht, n, _ in pins:
                    ing.push_batch(hp, ho, ht, pinned=True)
            ing.flush()
            rows = 0
            if ing_ex is not None:
                # N > 1: the step's result is the OWNERS' tables: window close over NVLink, then each rank reads
                # the rows it owns back to the host
                ing_ex.close_window(None)
                for w in ing_owned.list_windows():
                    rows += len(ing_owned.snapshot(w, cap=4096))
                ing_owned.reset_device()
                return rows
            for w in ing.list_windows():
                rows += len(ing.snapshot(w, cap=4096))
            return rows

        e2e_step()
        barrier()
        t0 = time.perf_counter()
        e2e_steps = max(1, min(args.steps, 3))
        rows = 0
        for _ in range(e2e_steps):
            rows = e2e_step()
        barrier()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = tt.item()
        k2 = ing.counters()
        assert k2["events_ok"] == events * (e2e_steps + 1), k2
        e2e = {"value": world * events * e2e_steps / dt, "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": rows * 48 + 8 * 40, "steps": e2e_steps,
               "bytes_per_event_on_the_link": h2d / events,
               "path": "bh_push_batch_compact (u16 lengths + i16 timestamp deltas rebuilt on the GPU) -> (pinned host -> HBM in "
                       "64 MiB ring slabs on 2 copy streams) -> k_process_raw -> bh_flush -> " + ("window close to the key owners over NVLink -> bh_snapshot of the owned windows"
                                                          if ing_ex is not None else "bh_snapshot of every window"),
               "pcie_gbs": world * h2d * e2e_steps / dt / 1e9,
               "pcie_peak_gbs_per_gpu": pcie_peak}
        if pcie_concurrent is None:
            e2e.update({"pcie_frac": e2e["pcie_gbs"] / pcie_peak,
                        "bound": "PCIe H2D ingest (measured pinned-copy peak of this GPU in pcie_peak_gbs_per_gpu)"})
        else:
            e2e.update({"pcie_concurrent_peak_gbs": pcie_concurrent, "pcie_concurrent_sum_of_rates_gbs": pcie_concurrent_sum,
                        "pcie_frac": e2e["pcie_gbs"] / pcie_concurrent,
                        "pcie_frac_of_n_x_solo": e2e["pcie_gbs"] / (world * pcie_peak),
                        "bound": "PCIe H2D ingest with all N GPUs copying at once (pcie_concurrent_peak_gbs = all bytes / time "
                                 "until the slowest GPU is done, measured; pcie_peak_gbs_per_gpu is one GPU copying alone)"})
        if ing_ex is not None:
            ing_ex.close(); ing_owned.close()
        if world == 1:
            variants0 = {}
            e2e_step(False)
            t0 = time.perf_counter()
            for _ in range(2):
                e2e_step(False)
            d = (time.perf_counter() - t0) / 2
            variants0["pinned_u32_offsets_u64_ts"] = {"value": events / d, "unit": UNIT, "pcie_gbs": h2d_plain / d / 1e9, "pcie_frac": h2d_plain / d / 1e9 / pcie_peak,
                                                      "bytes_per_event_on_the_link": h2d_plain / events,
                                                      "path": "bh_push_batch_pinned (round 1's e2e path: u32 offsets + u64 timestamps travel with the payload)"}
            # the drop-in ingest paths: what a Node Buffer goes through.  (a) bh_push_batch from ordinary (pageable)
            # host memory: the library copies into its pinned ring first; (b) bh_push, one call per message, as
            # addon/beholder_napi.c does for each rmsg.message.content (index.js:129) -- driven from C so that no
            # Python call overhead is measured (bh_debug_push_each: a loop over bh_push).
            hp0, ho0, ht0, n0, total0 = pins[0][3], pins[0][4], pins[0][5], pins[0][6], pins[0][7]
            pg = (np.array(hp0, copy=True), np.array(ho0, copy=True), np.array(ht0, copy=True))     # pageable copies
            b0 = total0 + (n0 + 1) * 4 + n0 * 8
            variants = variants0
            for name, fn in (("pageable_batch", lambda: ing.push_batch(*pg)),
                             ("per_message_push", lambda: L.check(L.lib.bh_debug_push_each(ing._ctx, pg[0].ctypes.data, pg[1].ctypes.data,
                                                                                          pg[2].ctypes.data, n0), ing._ctx))):
                fn(); ing.flush()
                t0 = time.perf_counter()
                reps = 3
                for _ in range(reps):
                    fn()
                ing.flush()
                d = (time.perf_counter() - t0) / reps
                variants[name] = {"value": n0 / d, "unit": UNIT, "events": n0, "pcie_gbs": b0 / d / 1e9, "pcie_frac": b0 / d / 1e9 / pcie_peak}
            # (c) the caller's own (ordinary) buffers page-locked once with bh_host_register, as a reused buffer pool
            # would be, then handed over without a copy
            for arr in pg:
                L.check(L.lib.bh_host_register(arr.ctypes.data, arr.nbytes))
            try:
                fn = lambda: ing.push_batch(*pg, pinned=True)   # noqa: E731
                fn(); ing.flush()
                t0 = time.perf_counter()
                for _ in range(3):
                    fn()
                ing.flush()
                d = (time.perf_counter() - t0) / 3
                variants["registered_batch"] = {"value": n0 / d, "unit": UNIT, "events": n0, "pcie_gbs": b0 / d / 1e9, "pcie_frac": b0 / d / 1e9 / pcie_peak,
                                                "path": "caller's buffers page-locked once (bh_host_register) -> bh_push_batch_pinned -> H2D in place -> kernel"}
            finally:
                ing.flush()
                for arr in pg:
                    L.lib.bh_host_unregister(arr.ctypes.data)
            variants["pageable_batch"]["limit"] = ("one pass over host DRAM: the copy into the pinned ring (8 threads) shares the host's memory bandwidth "
                                                   "with the DMA reads; this VM copies ~20 GB/s with any number of threads")
            variants["pageable_batch"]["path"] = "bh_push_batch (pageable host slab -> memcpy into the pinned ring on copy threads -> H2D -> kernel)"
            variants["per_message_push"]["path"] = "bh_push per message from C (what the N-API shim does per rmsg) -> pinned ring -> H2D -> kernel"
            e2e["variants"] = variants
        ing.close()
        for pb, ob, tb, *_ in pins:
            pb.free(); ob.free(); tb.free()
        for bufs, *_ in cpins:
            for b in bufs:
                b.free()

    if rank == 0:
        peak, peak_src = peaks()
        # our kernels per step: one fused launch per slab; with N > 1 also export (count, prefix, scatter) + merge
        launches_per_step = SLABS + (exch.kernel_launches_per_close if exch else 0)
        per_launch_s = kernel_ms * 1e-3 / (args.steps * SLABS)
        per_launch_bytes = sum(total + (n + 1) * 4 + n * 8 for *_, n, total in slabs) / SLABS
        achieved = per_launch_bytes / per_launch_s / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64 counts / f64 sum,min,max", "data": "synthetic", "config": workload_config(world),
            "roofline": {"bound": "hbm", "kernel": "bh::k_process_raw", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": ncu_traffic(), "peak_source": peak_src,
                         "algorithmic_bytes_per_event": BYTES_PER_EVENT, "events_per_launch": events // SLABS,
                         "avg_launch_ms": per_launch_s * 1e3},
            "gpu_launches": launches_per_step * args.steps,
            "exchange": (exchange if world > 1 else None),
            "exchange_check": (exchange_check if world > 1 else None),
            "clocks": clocks,
        }
        if e2e:
            line["e2e"] = e2e
        if ref_mode:
            ref_mode["frac_of_hbm_peak"] = ref_mode["achieved_gbs"] / peak
            line["reference_mode"] = ref_mode
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline_block(True)
    # ---- BASELINE.json configs 3, 4, 5 (extra keys of the same line; a failure is reported, not fatal) ----
    extra = {}
    # the N-GPU link peak the c5 leg compares with: known to every rank (rank 0 alone holds `e2e`)
    pcie_conc_all = (pcie_concurrent if (not args.no_e2e and world > 1 and pcie_concurrent) else 55.0 * world)
    if not args.no_extra:
        local.close(); local = None
        if exch is not None and hasattr(exch, "close"):
            exch.close()
        del slabs
        torch.cuda.empty_cache()
        peak_gbs = peaks()[0]
        legs = [("c3", lambda: c3_leg(args, local_rank, stream, peak_gbs)),
                ("c5", lambda: c5_leg(local_rank, stream, peak_gbs, e2e["pcie_peak_gbs_per_gpu"] if e2e else 55.0))] if world == 1 else \
               [("c4", lambda: c4_leg(args, world, rank, local_rank, stream, peak_gbs)),
                ("c5", lambda: c5_leg(local_rank, stream, peak_gbs, pcie_conc_all, world, rank))]
        # The headline above must never be lost to an extra leg.  One GPU: an exception is recorded in the leg's
        # place.  Several GPUs: a rank that failed would leave its peers waiting in a barrier for ever, so every
        # rank arms a watchdog: past the deadline rank 0 prints the line with the leg marked as timed out and all
        # ranks leave.
        def give_up():
            if rank == 0:
                line.update(extra)
                for name, _ in legs:
                    line.setdefault(name, {"error": "leg did not finish within its deadline"})
                print(json.dumps(line), flush=True)
            os._exit(0)
        dog = threading.Timer(float(os.environ.get("BEHOLDER_EXTRA_DEADLINE_S", "240")), give_up) if world > 1 else None
        if dog:
            dog.daemon = True; dog.start()
        for name, fn in legs:
            try:
                extra[name] = fn()
            except Exception as e:      # noqa: BLE001
                extra[name] = {"error": f"{type(e).__name__}: {e}"}
                if world > 1:
                    print(f"[bench] rank {rank}: {name} failed: {e}", file=sys.stderr, flush=True)
                    if dog:
                        dog.cancel()
                    time.sleep(5)       # peers may be inside the leg's barriers: let the watchdogs of the others end the run
                    give_up()
        if dog:
            dog.cancel()
    if rank == 0:
        line.update(extra)
        print(json.dumps(line))
    if local is not None:
        local.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

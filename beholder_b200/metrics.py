"""`/metrics` exposition of the device tables: Prometheus text format for the windowed
aggregates, including the log2 histograms in cumulative `_bucket{le=...}` form, so that scrapers of
the reference's `Prom.expose()` endpoint (index.js:28) keep working.  The reference itself only
exposes `beholder_progress_updates_total{status}` (handler.GpuCounter renders that one); the series
below belong to the windowed extension and have no reference counterpart.
"""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np

from . import _lib as L


def _num(v) -> str:
    """prom-client prints JS numbers: integers without a fraction, +Inf/-Inf/NaN spelled out"""
    f = float(v)
    if f != f:
        return "NaN"
    if f in (float("inf"), float("-inf")):
        return "+Inf" if f > 0 else "-Inf"
    if f == int(f) and abs(f) < 1e21:
        return str(int(f))
    return repr(f)


def bucket_upper_bounds(buckets: int = 64, e0: int = 991):
    """`le` of each histogram bucket: bucket 0 holds v <= 0; bucket k >= 1 holds values whose biased
    exponent is e0 + k (bucket 1 also everything smaller, the last bucket everything larger)."""
    le = ["0"]
    for k in range(1, buckets - 1):
        le.append(_num(2.0 ** (k + e0 - 1022)))
    le.append("+Inf")
    return le


def render_windowed(records: np.ndarray, hist: Optional[np.ndarray] = None, *, name: str = "beholder_progress",
                    ops: int = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX, hist_e0: int = 991,
                    key_label: Callable[[int], str] = lambda kh: f"{kh:016x}") -> str:
    order = np.lexsort((records["key_hash"], records["window_id"]))
    out = []

    def series(suffix, typ, help_, col):
        out.append(f"# HELP {name}_{suffix} {help_}")
        out.append(f"# TYPE {name}_{suffix} {typ}")
        for i in order:
            r = records[i]
            out.append(f'{name}_{suffix}{{key="{key_label(int(r["key_hash"]))}",window="{int(r["window_id"])}"}} {_num(r[col])}')

    if hist is not None and (ops & L.OP_HIST):
        B = hist.shape[1]
        le = bucket_upper_bounds(B, hist_e0)
        out.append(f"# HELP {name}_value log2-bucket histogram of the reduced value per key and window")
        out.append(f"# TYPE {name}_value histogram")
        for i in order:
            r = records[i]
            lab = f'key="{key_label(int(r["key_hash"]))}",window="{int(r["window_id"])}"'
            cum = np.cumsum(hist[i].astype(np.uint64))
            for k in range(B):
                out.append(f'{name}_value_bucket{{{lab},le="{le[k]}"}} {int(cum[k])}')
            out.append(f"{name}_value_sum{{{lab}}} {_num(r['sum'])}")
            out.append(f"{name}_value_count{{{lab}}} {int(r['count'])}")
    series("events_total", "counter", "events aggregated per key and window", "count")
    if ops & L.OP_SUM:
        series("sum", "gauge", "sum of the reduced value per key and window", "sum")
    if ops & L.OP_MIN:
        series("min", "gauge", "minimum of the reduced value per key and window", "min")
    if ops & L.OP_MAX:
        series("max", "gauge", "maximum of the reduced value per key and window", "max")
    return "\n".join(out) + "\n"


def render_aggregator(agg, **kw) -> str:
    """the exposition of everything an Aggregator currently holds"""
    with_hist = bool(agg.ops & L.OP_HIST)
    snap = agg.snapshot_all(with_hist=with_hist)
    recs, hist = snap if with_hist else (snap, None)
    return render_windowed(recs, hist, ops=agg.ops, hist_e0=agg.cfg.hist_e0, **kw)

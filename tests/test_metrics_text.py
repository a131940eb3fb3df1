"""/metrics exposition of windowed aggregates: format and histogram bucket semantics,
checked on the oracle's rows (no GPU)."""
import re

import numpy as np

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.metrics import bucket_upper_bounds, render_windowed
from beholder_b200.synth import SynthSpec, generate_host
from oracle import oracle as O
from helpers import ALL_OPS, SEED


def test_bucket_bounds_match_the_bucket_function():
    le = bucket_upper_bounds(64, 991)
    assert le[0] == "0" and le[-1] == "+Inf" and len(le) == 64
    for k in range(1, 63):
        ub = float(le[k])
        assert O.lib.bho_hist_bucket(np.nextafter(ub, 0), 64, 991) == k      # just below the bound: bucket k
        assert O.lib.bho_hist_bucket(ub, 64, 991) == k + 1                   # the bound itself: next bucket
    assert le[32] == "2"                                                     # bucket 32 is [1, 2)


def test_render_is_valid_prometheus_text_with_cumulative_buckets():
    spec = SynthSpec(n_keys=12, rate_per_ms=1, value_f64=True, value_signed=True)
    slab, off, ts = generate_host(spec, 0, 3000)
    schema = bh.telemetry_progress(L.K_DOUBLE).with_roles(key=("mediaId", "status"))
    ops = ALL_OPS | L.OP_HIST
    a = O.Agg(schema, ops, 10_000, SEED)
    a.push_raw(slab, off, ts)
    recs, hist = a.snapshot(with_hist=True)
    text = render_windowed(recs.view(bh.RECORD_DTYPE), hist, ops=ops)
    lines = text.splitlines()
    pat = re.compile(r'^[a-z_]+(\{[a-z]+="[^"]*"(,[a-z]+="[^"]*")*\})? (-?[0-9.e+-]+|\+Inf|-Inf|NaN)$')
    for ln in lines:
        assert ln.startswith("# HELP ") or ln.startswith("# TYPE ") or pat.match(ln), ln
    # per series: buckets are cumulative, +Inf equals _count equals the row's count
    inf = [int(l.rsplit(" ", 1)[1]) for l in lines if 'le="+Inf"' in l]
    cnt = [int(l.rsplit(" ", 1)[1]) for l in lines if l.startswith("beholder_progress_value_count")]
    assert inf == cnt and sum(cnt) == int(recs["count"].sum()) == 3000
    first = [int(l.rsplit(" ", 1)[1]) for l in lines if l.startswith("beholder_progress_value_bucket")][:64]
    assert first == sorted(first)
    assert sum(1 for l in lines if l.startswith("beholder_progress_events_total{")) == len(recs)

"""One process, two group members on ONE GPU (so that ncu sees every kernel): a C4-like window close at reduced
scale -- per-kernel times of fold, export-to-owner, merge, recycle.
    ncu --metrics gpu__time_duration.sum --csv --log-file out.csv python tools/probe/c4_close_profile.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.dist import Group
from beholder_b200.synth import SynthSpec, generate_device

KEYS, N = 5_000_000, 12_500_000
ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX
spec = SynthSpec(n_keys=KEYS, rate_per_ms=5000, jitter_ms=2000, payload_bytes=64)
g = Group([0, 0], owner_stride=int(KEYS * 2), owned_table_capacity=1 << 22, mode=L.MODE_WINDOWED, ops=ops, window_ms=10_000,
          table_capacity=1 << 23, window_ring=4)
slabs = [generate_device(spec, i * N, N) for i in range(4)]
torch.cuda.synchronize()
for rnd in range(2):
    for r in range(2):
        for k in range(2):
            p, o, t, _ = slabs[r * 2 + k]
            g.local(r).process_device(p, o, t, N)
    g.close_window()
g.sync()
rows = sum(len(g.owned(r).snapshot_all()) for r in range(2))
print("owned rows", rows)
g.close()

"""Host-side ingest paths on one GPU: pinned batch, pageable batch (library copies into its ring), per-message push."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_device
n = 12_500_000
p, o, t, total = generate_device(SynthSpec(n_keys=1000, payload_bytes=64), 0, n)
hp = p[: total + 16].cpu().numpy(); ho = o.cpu().numpy().view(np.uint32); ht = t.cpu().numpy().view(np.uint64)
ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX
b = total + 12 * n
# raw host memcpy bandwidth, 1 thread and numpy
dst = np.empty_like(hp)
t0 = time.perf_counter(); np.copyto(dst, hp); t1 = time.perf_counter()
print("numpy copy GB/s", hp.nbytes / (t1 - t0) / 1e9, "cpus", os.cpu_count())
for slab_mb in (64, 256):
    g = bh.Aggregator(mode=L.MODE_WINDOWED, ops=ops, table_capacity=2048, window_ring=32, slab_bytes=slab_mb << 20, slab_events=(slab_mb << 20) // 64, n_slabs=4)
    g.push_batch(hp, ho, ht); g.flush()
    t0 = time.perf_counter()
    for _ in range(3): g.push_batch(hp, ho, ht)
    g.flush(); dt = (time.perf_counter() - t0) / 3
    print(json.dumps({"slab_mb": slab_mb, "threads": os.environ.get("BEHOLDER_COPY_THREADS"), "pageable_GBps": b / dt / 1e9, "ev_s": n / dt}))
    g.close()

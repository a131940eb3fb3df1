"""N-rank check (torchrun): dist.MemberExchange (the library's own multi-process exchange: CUDA IPC receive areas,
export kernel stores rows into the owners' buffers over NVLink, flag barrier in peer memory, merge) gives every
owner exactly the rows the CPU oracle computes for its keys on the WHOLE stream; also times a window close."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tests"))
import numpy as np, torch, torch.distributed as dist
import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.dist import MemberExchange, owner_of
from beholder_b200.synth import SynthSpec, generate_device
from oracle import oracle as O

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
os.environ.setdefault("NCCL_DEBUG", "WARN")
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
st = torch.cuda.Stream(); torch.cuda.set_stream(st); s = st.cuda_stream
HIST = bool(int(os.environ.get("PEER_HIST", "0")))
ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX | (L.OP_HIST if HIST else 0)
KEY = ("mediaId", "status")
schema = bh.telemetry_progress(L.K_DOUBLE if HIST else L.K_INT32)
mk = lambda: bh.Aggregator(schema, mode=L.MODE_WINDOWED, key=KEY, ops=ops, window_ms=10_000, table_capacity=1 << 14, window_ring=8, device=lr)
n, rounds = 400_000, 3
spec = SynthSpec(n_keys=4000, rate_per_ms=50, value_f64=HIST, value_exp_lo=-20, value_exp_span=40, malformed_ppm=1000)
local, owned = mk(), mk()
ex = MemberExchange(local, owned, world, rank, owner_stride=40_000, device=lr, hist_buckets=64 if HIST else 0)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
tot = 0.0
for k in range(rounds):
    p, o, t, _ = generate_device(spec, (k * world + rank) * n, n, device=lr)
    local.process_device(p, o, t, n, stream=s)
    e0.record(); ex.close_window(None, s); e1.record(); torch.cuda.synchronize()
    tot += e0.elapsed_time(e1)
ex.check_no_overflow()
owned.sync()
got = owned.snapshot_all(with_hist=HIST)
# the oracle on the whole stream (every rank's share), reduced to the keys this rank owns
a = O.Agg(schema.with_roles(key=KEY), ops, 10_000, 0xBE401DE5)
for g in range(rounds * world):
    p, o, t, _ = generate_device(spec, g * n, n, device=lr)
    a.push_raw(p.cpu().numpy(), o.cpu().numpy().view(np.uint32), t.cpu().numpy().view(np.uint64))
from helpers import assert_records_equal
if HIST:
    want, wh = a.snapshot(with_hist=True)
    mine = owner_of(want["key_hash"], world) == rank
    assert_records_equal(got[0], want[mine], ops, got[1], wh[mine])
    rows = len(got[0])
else:
    want = a.snapshot()
    assert_records_equal(got, want[owner_of(want["key_hash"], world) == rank], ops)
    rows = len(got)
print(json.dumps({"rank": rank, "rows": rows, "oracle_equal": True, "close_ms": tot / rounds}), flush=True)
dist.barrier(); ex.close(); local.close(); owned.close(); dist.destroy_process_group()

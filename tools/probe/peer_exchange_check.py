"""N-rank check (torchrun): PeerWindowExchange gives the owners the same tables as the NCCL WindowExchange,
and how long a window close takes with each."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch, torch.distributed as dist
import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.dist import WindowExchange, PeerWindowExchange
from beholder_b200.synth import SynthSpec, generate_device

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
os.environ.setdefault("NCCL_DEBUG", "WARN")
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
HIST = bool(int(os.environ.get("PEER_HIST", "0")))
ops = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX | (L.OP_HIST if HIST else 0)
mk = lambda: bh.Aggregator(mode=L.MODE_WINDOWED, ops=ops, window_ms=10_000, table_capacity=2048, window_ring=32, device=lr)
n = 2_000_000
spec = SynthSpec(n_keys=1000, rate_per_ms=500)
p, o, t, _ = generate_device(spec, rank * n, n, device=lr)
stride = 2 * 1000 * 40 // world


class HistNccl(WindowExchange):
    """the NCCL exchange with histogram rows (two all-to-alls), the reference point for PEER_HIST=1"""
    def close_all_windows(self, reset_local=True):
        from beholder_b200._lib import BH_ALL_WINDOWS
        from beholder_b200.dist import all_to_all_records
        s_ = torch.cuda.current_stream().cuda_stream
        if not hasattr(self, "send_hist"):
            self.send_hist = torch.zeros(self.world, self.owner_stride, 64, dtype=torch.int32, device=self.send.device)
        self.send.zero_()
        self.local.export_window_device(BH_ALL_WINDOWS, self.world, self.send, self.send_hist, self.world * self.owner_stride,
                                        self.counts, owner_stride=self.owner_stride, stream=s_)
        recv = all_to_all_records(self.send)
        rh = torch.empty_like(self.send_hist)
        dist.all_to_all_single(rh.view(-1), self.send_hist.view(-1))
        self.owned.merge_records_device(recv, rh, self.world * self.owner_stride, stream=s_)
        if reset_local:
            self.local.reset_device(stream=s_)


res = {}
for name, cls in (("nccl", WindowExchange), ("peer", PeerWindowExchange)):
    local, owned = mk(), mk()
    if HIST and cls is WindowExchange:
        ex = HistNccl(local, owned, world, rank, owner_stride=stride, device=lr)
    elif HIST:
        ex = cls(local, owned, world, rank, owner_stride=stride, device=lr, hist_buckets=64)
    else:
        ex = cls(local, owned, world, rank, owner_stride=stride, device=lr)
    for _ in range(3):
        local.process_device(p, o, t, n, stream=st.cuda_stream)
        if os.environ.get("PEER_DEBUG"):
            torch.cuda.synchronize(); print(rank, name, "processed", local.counters(check=False), flush=True)
        ex.close_all_windows()
        if os.environ.get("PEER_DEBUG"):
            torch.cuda.synchronize(); print(rank, name, "closed", flush=True)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for _ in range(20):
        local.process_device(p, o, t, n, stream=st.cuda_stream)
        e0.record(); ex.close_all_windows(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    owned.sync()
    if HIST:
        recs, hist = owned.snapshot_all(with_hist=True)
        order = np.lexsort((recs["key_hash"], recs["window_id"]))
        recs, hist = recs[order], hist[order]
    else:
        recs = owned.snapshot_all()
        recs = recs[np.lexsort((recs["key_hash"], recs["window_id"]))]
        hist = None
    res[name] = (recs, tot / 20, hist)
    ex.check_no_overflow()
    local.close(); owned.close()
a, b = res["nccl"][0], res["peer"][0]
assert len(a) == len(b) and len(a) > 0, (len(a), len(b))
for f in ("key_hash", "window_id", "count", "min", "max"):
    assert np.array_equal(a[f], b[f]), f
assert np.allclose(a["sum"], b["sum"], rtol=1e-9)
if HIST:
    assert np.array_equal(res["nccl"][2], res["peer"][2]) and res["peer"][2].sum() == a["count"].sum()
print(json.dumps({"rank": rank, "rows": len(a), "close_ms_nccl": res["nccl"][1], "close_ms_peer": res["peer"][1]}), flush=True)
dist.barrier(); dist.destroy_process_group()

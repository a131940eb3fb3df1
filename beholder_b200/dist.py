"""Multi-GPU window close: key-hash sharded merge of per-window partial aggregates.

One process per GPU (torch.distributed; NCCL over NVLink on the GPU box, gloo in the CPU tests).
Events are spread over the ranks by arrival; every rank aggregates whatever keys it sees.  At
window close each rank compacts its rows grouped by OWNER (owner = top log2(world) bits of the
key hash), the ranks exchange rows with one fixed-size all-to-all, and each owner merges what it
received into its own table: count/sum/hist add, min/max.  Integer results do not depend on
the number of ranks; f64 sums differ by reduction order only.  The reference has no counterpart
(it is a single Node process).
"""
from __future__ import annotations

from typing import Optional

import numpy as np

RECORD_WORDS = 6   # bh_record = 6 x 8 bytes


def owner_of(key_hash: np.ndarray, world: int) -> np.ndarray:
    """owner rank of each key: its top log2(world) bits (world is a power of two)"""
    lg = int(world).bit_length() - 1
    assert 1 << lg == world, "world size must be a power of two"
    if lg == 0:
        return np.zeros(len(key_hash), dtype=np.int64)
    return (key_hash.astype(np.uint64) >> np.uint64(64 - lg)).astype(np.int64)


def partition_by_owner(records: np.ndarray, world: int, owner_stride: int) -> np.ndarray:
    """Host mirror of bh_export_window_device(owner_stride > 0): a [world, owner_stride] array of
    records, zero rows (count 0) as padding.  Raises if an owner has more rows than the stride."""
    out = np.zeros((world, owner_stride), dtype=records.dtype)
    own = owner_of(records["key_hash"], world)
    for o in range(world):
        rows = records[own == o]
        if len(rows) > owner_stride:
            raise ValueError(f"owner {o} has {len(rows)} rows > owner_stride {owner_stride}")
        out[o, : len(rows)] = rows
    return out


def all_to_all_records(send, group=None):
    """send: torch tensor [world, owner_stride, 6] int64 (device or CPU).  Row block o goes to rank o;
    returns the blocks received from every rank, same shape."""
    import torch
    import torch.distributed as dist
    recv = torch.empty_like(send)
    if dist.get_world_size(group) == 1:
        recv.copy_(send)
        return recv
    dist.all_to_all_single(recv.view(-1), send.view(-1), group=group)
    return recv


def _gcheck(rc: int) -> None:
    from . import _lib as L
    if rc != L.BH_OK:
        raise L.BeholderError(rc, (L.lib.bh_group_last_error() or b"").decode())


class Group:
    """bh_create_group: one process drives several GPUs (the single process of index.js:23-48 is what it serves).
    `local(r)` takes GPU r's share of the events, `owned(r)` holds the merged rows of the keys GPU r owns,
    `close_window` runs export-to-peers -> barrier -> merge -> recycle on every GPU.  No PyTorch, no NCCL."""

    def __init__(self, devices, owner_stride: int, owned_table_capacity: int = 0, **agg_kwargs):
        import ctypes as C
        from . import _lib as L
        from .engine import Aggregator
        lc = Aggregator.config(**agg_kwargs)
        okw = dict(agg_kwargs)
        if owned_table_capacity:
            okw["table_capacity"] = owned_table_capacity
        oc = Aggregator.config(**okw)
        devs = (C.c_int * len(devices))(*devices)
        self._g = C.c_void_p()
        _gcheck(L.lib.bh_create_group(C.byref(lc), C.byref(oc), devs, len(devices), owner_stride, C.byref(self._g)))
        self.world = len(devices)
        self._local = [Aggregator(_handle=L.lib.bh_group_local(self._g, r), device=devices[r], **agg_kwargs) for r in range(self.world)]
        self._owned = [Aggregator(_handle=L.lib.bh_group_owned(self._g, r), device=devices[r], **okw) for r in range(self.world)]

    def local(self, rank: int):
        return self._local[rank]

    def owned(self, rank: int):
        return self._owned[rank]

    def close_window(self, window_id: Optional[int] = None) -> None:
        from . import _lib as L
        _gcheck(L.lib.bh_group_close_window(self._g, L.BH_ALL_WINDOWS if window_id is None else window_id))

    def sync(self) -> None:
        from . import _lib as L
        _gcheck(L.lib.bh_group_sync(self._g))

    def close(self) -> None:
        from . import _lib as L
        if self._g:
            L.lib.bh_group_destroy(self._g)
            self._g = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class MemberExchange:
    """One rank of a multi-process group (one process per GPU, as torchrun starts them): bh_member_* of the
    library.  The only thing torch.distributed does here is carry the 64-byte CUDA IPC handles of the receive
    areas between the ranks once, at construction (`all_gather_object`); every window close afterwards is the
    library's own kernels: export rows into the owners' buffers over NVLink, a flag barrier in peer memory,
    merge.  One barrier per close (receive buffers alternate), nothing is cleared."""

    def __init__(self, local, owned, world: int, rank: int, owner_stride: int, device: int = 0, hist_buckets: int = 0,
                 group=None):
        import ctypes as C
        import torch.distributed as dist
        from . import _lib as L
        self.local, self.owned, self.world, self.rank, self.owner_stride = local, owned, world, rank, owner_stride
        self._m = C.c_void_p()
        _gcheck(L.lib.bh_member_create(local._ctx, owned._ctx, device, rank, world, owner_stride, hist_buckets, C.byref(self._m)))
        h = C.create_string_buffer(64)
        _gcheck(L.lib.bh_member_ipc_handle(self._m, h))
        handles = [None] * world
        dist.all_gather_object(handles, h.raw, group=group)
        blob = C.create_string_buffer(b"".join(handles), 64 * world)
        rc = L.lib.bh_member_connect_ipc(self._m, blob)
        err = (L.lib.bh_group_last_error() or b"").decode()
        # every rank learns whether every rank has mapped every area (this also orders the mappings before the
        # first store); a failure anywhere raises everywhere, so the callers can fall back together
        oks = [None] * world
        dist.all_gather_object(oks, rc == L.BH_OK, group=group)
        if not all(oks):
            self.close()
            raise L.BeholderError(rc if rc != L.BH_OK else L.BH_ERR_CUDA, err or "a peer rank could not map the receive areas")
        self.kernel_launches_per_close = 3 + 2 + 1 + 2   # export (prefix, scatter, counts), post counts + barrier, merge, close

    def close_window(self, window_id: Optional[int] = None, stream: int = 0) -> None:
        from . import _lib as L
        _gcheck(L.lib.bh_member_close_window(self._m, L.BH_ALL_WINDOWS if window_id is None else window_id, stream or None))

    def close_all_windows(self, reset_local: bool = True):
        import torch
        self.close_window(None, torch.cuda.current_stream().cuda_stream or 1)     # 1 = cudaStreamLegacy, never "the context's stream"

    def check_no_overflow(self):
        import ctypes as C
        from . import _lib as L
        cnt = (C.c_uint64 * self.world)()
        _gcheck(L.lib.bh_member_check(self._m, cnt))
        return list(cnt)

    def close(self) -> None:
        from . import _lib as L
        if self._m:
            L.lib.bh_member_destroy(self._m)
            self._m = None


class WindowExchange:
    """GPU side: export (grouped by owner) -> all-to-all over NCCL -> merge into the owner's table.

    `local` aggregates this rank's share of the events; `owned` holds the merged rows of the keys
    this rank owns.  All steps are enqueued on torch's current stream; nothing synchronises the host.
    """

    def __init__(self, local, owned, world: int, rank: int, owner_stride: int, device: int = 0):
        import torch
        self.local, self.owned, self.world, self.rank = local, owned, world, rank
        self.owner_stride = owner_stride
        dev = torch.device("cuda", device)
        self.send = torch.zeros(world, owner_stride, RECORD_WORDS, dtype=torch.int64, device=dev)
        self.counts = torch.zeros(world, dtype=torch.int64, device=dev)
        self.kernel_launches_per_close = 3 + 1   # export (count, prefix, scatter) + merge

    def close_all_windows(self, reset_local: bool = True):
        import torch
        from ._lib import BH_ALL_WINDOWS
        # torch's default stream has handle 0, which the C-ABI reads as "the context's own stream": name the
        # legacy default stream explicitly (cudaStreamLegacy = 1) so that every step really is on torch's stream
        st = torch.cuda.current_stream().cuda_stream or 1
        self.send.zero_()
        self.local.export_window_device(BH_ALL_WINDOWS, self.world, self.send, None, self.world * self.owner_stride,
                                        self.counts, owner_stride=self.owner_stride, stream=st)
        recv = all_to_all_records(self.send)
        self.owned.merge_records_device(recv, None, self.world * self.owner_stride, stream=st)
        if reset_local:
            self.local.reset_device(stream=st)
        return recv

    def check_no_overflow(self):
        """host check (synchronises): no owner block was cut"""
        c = self.counts.cpu().numpy()
        if (c > self.owner_stride).any():
            raise RuntimeError(f"owner_stride {self.owner_stride} too small: rows per owner {c.tolist()}")


class PeerWindowExchange(WindowExchange):
    """Window close over peer memory instead of NCCL: every rank's receive buffer is a torch
    symmetric-memory tensor (mapped into all ranks of the node, NVLink / NVSwitch), and the export kernel
    of rank r stores owner o's rows (and histograms, with `hist_buckets`) straight into block r of GPU o's
    buffer (`bh_export_window_peers`): the compaction and the exchange are one kernel, there is no send
    staging buffer and no all-to-all.  Two stream-ordered cross-GPU barriers per close frame the stores:
    one after them (owners may merge), one after the owner has merged and re-zeroed its buffer (peers may
    store again).  Same results as WindowExchange.  Needs all ranks on one node."""

    def __init__(self, local, owned, world: int, rank: int, owner_stride: int, device: int = 0, group=None,
                 hist_buckets: int = 0):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.local, self.owned, self.world, self.rank = local, owned, world, rank
        self.owner_stride = owner_stride
        dev = torch.device("cuda", device)
        group = group or dist.group.WORLD
        self.recv = symm.empty((world, owner_stride, RECORD_WORDS), dtype=torch.int64, device=dev)
        self.recv.zero_()
        self.hdl = symm.rendezvous(self.recv, group)
        self.peer_bases = torch.tensor([int(p) for p in self.hdl.buffer_ptrs], dtype=torch.int64, device=dev)
        self.recv_hist = self.peer_hist_bases = None
        if hist_buckets:
            self.recv_hist = symm.empty((world, owner_stride, hist_buckets), dtype=torch.int32, device=dev)
            self.recv_hist.zero_()
            self.hdl_hist = symm.rendezvous(self.recv_hist, group)
            self.peer_hist_bases = torch.tensor([int(p) for p in self.hdl_hist.buffer_ptrs], dtype=torch.int64, device=dev)
        self.counts = torch.zeros(world, dtype=torch.int64, device=dev)
        self.kernel_launches_per_close = 3 + 1   # export (count, prefix, scatter-to-peers) + merge
        torch.cuda.synchronize(dev)
        self.hdl.barrier(channel=0)              # every buffer is zero before anyone stores

    def close_all_windows(self, reset_local: bool = True):
        import torch
        from ._lib import BH_ALL_WINDOWS
        st = torch.cuda.current_stream().cuda_stream or 1      # see WindowExchange
        self.local.export_window_peers(BH_ALL_WINDOWS, self.world, self.peer_bases, self.rank, self.owner_stride,
                                       self.counts, d_peer_hist_bases=self.peer_hist_bases, stream=st)
        self.hdl.barrier(channel=0)              # all rows have arrived
        self.owned.merge_records_device(self.recv, self.recv_hist, self.world * self.owner_stride, stream=st)
        self.recv.zero_()                        # (histogram rows of padding records are never read)
        if reset_local:
            self.local.reset_device(stream=st)
        self.hdl.barrier(channel=1)              # nobody stores into a buffer that is still being merged / cleared
        return self.recv

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
        st = torch.cuda.current_stream().cuda_stream
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
        st = torch.cuda.current_stream().cuda_stream
        self.local.export_window_peers(BH_ALL_WINDOWS, self.world, self.peer_bases, self.rank, self.owner_stride,
                                       self.counts, d_peer_hist_bases=self.peer_hist_bases, stream=st)
        self.hdl.barrier(channel=0)              # all rows have arrived
        self.owned.merge_records_device(self.recv, self.recv_hist, self.world * self.owner_stride, stream=st)
        self.recv.zero_()                        # (histogram rows of padding records are never read)
        if reset_local:
            self.local.reset_device(stream=st)
        self.hdl.barrier(channel=1)              # nobody stores into a buffer that is still being merged / cleared
        return self.recv

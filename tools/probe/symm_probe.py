"""Does torch's symmetric memory (NVLink peer buffers + stream-ordered barrier) work on this box?"""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
os.environ.setdefault("NCCL_DEBUG", "WARN")
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
buf = symm.empty((world, 16), dtype=torch.int64, device=torch.device("cuda", lr))
buf.zero_()
hdl = symm.rendezvous(buf, dist.group.WORLD)
print(rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "dev", hex(hdl.buffer_ptrs_dev), "multicast", hdl.has_multicast_support, flush=True)
hdl.barrier(channel=0)
for peer in range(world):
    t = hdl.get_buffer(peer, (world, 16), torch.int64)
    t[rank].fill_(100 * rank + peer)
hdl.barrier(channel=0)
torch.cuda.synchronize()
print(rank, buf[:, 0].tolist(), flush=True)
assert buf[:, 0].tolist() == [100 * r + rank for r in range(world)]
dist.destroy_process_group()

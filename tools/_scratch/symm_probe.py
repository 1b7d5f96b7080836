import os, torch, torch.distributed as dist
rank=int(os.environ["RANK"]); world=int(os.environ["WORLD_SIZE"]); local=int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import torch.distributed._symmetric_memory as symm_mem
t = symm_mem.empty((world, 1024), dtype=torch.int64, device=torch.device("cuda", local))
t.zero_()
h = symm_mem.rendezvous(t, dist.group.WORLD.group_name if hasattr(dist.group.WORLD,'group_name') else dist.group.WORLD)
print(rank, "ptrs", [hex(p) for p in h.buffer_ptrs], "rank", h.rank, "world", h.world_size, flush=True)
# write my row into every peer through get_buffer
for p in range(world):
    pb = h.get_buffer(p, (world, 1024), torch.int64)
    pb[rank].fill_(100 + rank)
h.barrier()
torch.cuda.synchronize()
print(rank, "rows", t[:, 0].tolist(), flush=True)
dist.barrier()
dist.destroy_process_group()

"""Probe torch symmetric memory on this box (development): peer pointers, multicast support."""
import os, torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
import torch.distributed._symmetric_memory as symm
t = symm.empty(1 << 20, dtype=torch.float32, device=f"cuda:{rank}")
h = symm.rendezvous(t, dist.group.WORLD)
t.fill_(rank + 1)
h.barrier()
peer = h.get_buffer(1 - rank if world == 2 else (rank + 1) % world, (16,), torch.float32)
print(rank, "peer buffer value", peer[:2].tolist(), "ptrs", [hex(p) for p in h.buffer_ptrs][:4], "multicast_ptr", hex(getattr(h, "multicast_ptr", 0) or 0), flush=True)
h.barrier()
dist.destroy_process_group()

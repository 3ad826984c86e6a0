#!/usr/bin/env python
"""Does this box give torch symmetric memory a multicast (NVLS) mapping?  torchrun, one rank per GPU."""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
t = symm_mem.empty((1 << 20,), dtype=torch.uint8, device=dev)
h = symm_mem.rendezvous(t, dist.group.WORLD)
try:
    sup = type(h).has_multicast_support(torch.device("cuda").type if False else dev.type, local)
except Exception as e:  # signature differs between torch versions
    sup = f"query failed: {e}"
print(f"rank {rank}/{world}: multicast_ptr={int(h.multicast_ptr):#x} has_multicast_support={sup} buffer_ptrs={[hex(int(p)) for p in h.buffer_ptrs][:2]}", flush=True)
dist.barrier()
torch.cuda.synchronize()
os._exit(0)

import os, sys
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"])); dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
t = symm_mem.empty((1024, 1024), dtype=torch.bfloat16, device=dev)
h = symm_mem.rendezvous(t, dist.group.WORLD)
if rank == 0:
    print("multicast_ptr", hex(h.multicast_ptr) if h.multicast_ptr else 0, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "backend", symm_mem.get_backend(dev))
    try:
        from torch._C._distributed_c10d import _SymmetricMemory
        print("has_multicast_support", _SymmetricMemory.has_multicast_support(torch.device("cuda").type if False else "cuda", dev.index))
    except Exception as e: print("hms err", e)
dist.destroy_process_group()

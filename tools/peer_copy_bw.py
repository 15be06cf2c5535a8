#!/usr/bin/env python
"""torchrun target: copy-engine bandwidth of peer copies into symmetric memory: one 14.7 MB panel (2048 x 7168 bytes)
sent as 1 / 2 / 4 / 8 concurrent pieces on as many streams, rank r -> rank r+1, every rank at once."""
import json, os, sys
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
nbytes = 2048 * 7168
buf = symm_mem.empty((2 * nbytes,), dtype=torch.uint8, device=dev)
hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
peer = hdl.get_buffer((rank + 1) % world, (2 * nbytes,), torch.uint8, 0)
src = buf[:nbytes]
dst = peer[nbytes:]
streams = [torch.cuda.Stream(dev) for _ in range(8)]
res = {"world": world, "bytes": nbytes}
for pieces in (1, 2, 4, 8):
    step = nbytes // pieces
    def send():
        for i in range(pieces):
            with torch.cuda.stream(streams[i]):
                dst[i * step:(i + 1) * step].copy_(src[i * step:(i + 1) * step], non_blocking=True)
    for _ in range(3):
        send()
    torch.cuda.synchronize(); dist.barrier()
    ts = []
    for _ in range(10):
        e0 = torch.cuda.Event(enable_timing=True); e1s = [torch.cuda.Event(enable_timing=True) for _ in range(pieces)]
        torch.cuda.synchronize(); dist.barrier()
        e0.record()
        for s in streams[:pieces]:
            s.wait_event(e0)
        send()
        for i in range(pieces):
            e1s[i].record(streams[i])
        torch.cuda.synchronize()
        ts.append(max(e0.elapsed_time(e) for e in e1s))
    t = sorted(ts)[len(ts) // 2]
    res[f"pieces{pieces}_us"] = round(t * 1e3, 1)
    res[f"pieces{pieces}_GBps"] = round(nbytes / t / 1e6, 1)
if rank == 0:
    print(json.dumps(res), flush=True)
dist.destroy_process_group()

#!/usr/bin/env python
"""torchrun target (2+ GPUs): NVLink peer-write bandwidth as a function of the contiguous run length."""
import ctypes, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
# laboratory tool: needs the diagnostics build (make -C sgl-kernel-xpu_b200/csrc diag)
os.environ.setdefault("SGLK_LIBRARY", os.path.join(ROOT, "sgl-kernel-xpu_b200", "lib", "libsglk_diag.so"))
from sgl_kernel import _lib

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
rows, pitch = 16384, 36864
buf = symm_mem.empty((rows, pitch), dtype=torch.uint8, device=dev)
hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
lib = ctypes.CDLL(_lib.LIB_PATH)
f = lib.sglk_debug_peer_write
f.restype = ctypes.c_int
f.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
peer = int(hdl.buffer_ptrs[(rank + 1) % world])
local = int(hdl.buffer_ptrs[rank])
width = 4608  # bytes per row written (one 2304-column bf16 panel)
res = {}
for name, ptr in (("peer", peer), ("local", local)):
    for seg in (16, 32, 64, 128, 256, 512):
        st = torch.cuda.current_stream().cuda_stream
        f(ptr, pitch, rows, width, seg, 1, st)
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); rc = f(ptr, pitch, rows, width, seg, 20, st); e1.record(); torch.cuda.synchronize()
        assert rc == 0
        ms = e0.elapsed_time(e1) / 20
        res[f"{name}_seg{seg}"] = round(rows * width / ms / 1e6, 1)
        dist.barrier()
if rank == 0:
    print(json.dumps(res))
dist.destroy_process_group()

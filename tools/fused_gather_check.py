#!/usr/bin/env python
"""torchrun target: fused GEMM->all-gather vs NCCL all-gather (bitwise), timing of both."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
import torch.distributed as dist
import sgl_kernel
from sgl_kernel.sharded import ColumnShardedFp8Linear, shard_bounds

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
M, N, K = (int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (16384, 18432, 7168)))
lo, hi = shard_bounds(N, world, rank)
g = torch.Generator(device=dev).manual_seed(1234)
x = torch.randn(M, K, generator=g, device=dev).to(torch.bfloat16)
gw = torch.Generator(device=dev).manual_seed(77 + rank)
wq = ((torch.rand(hi - lo, K, generator=gw, device=dev) - 0.5) * 896).to(torch.float8_e4m3fn)
ws = (torch.rand(-((hi - lo) // -128), K // 128, generator=gw, device=dev) + 0.5) * 1e-3
layer = ColumnShardedFp8Linear(sgl_kernel.Fp8BlockLinear(wq, ws), N)
ref = layer(x, gather_output=True)
layer.enable_fused_gather(M)
out = layer.forward_fused_gather(x)
torch.cuda.synchronize(); dist.barrier()
same = bool(torch.equal(out, ref))
def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
out_mc = layer.forward_fused_gather(x, multicast=True).clone()
torch.cuda.synchronize(); dist.barrier()
same_mc = bool(torch.equal(out_mc, ref))
t_fused = timeit(lambda: layer.forward_fused_gather(x))
t_mc = timeit(lambda: layer.forward_fused_gather(x, multicast=True))
t_nccl = timeit(lambda: layer(x, gather_output=True))
t_local = timeit(lambda: layer(x, gather_output=False))
if rank == 0:
    print(json.dumps({"world": world, "M": M, "N": N, "K": K, "bitwise_equal_to_nccl_path": same, "ms_fused": t_fused, "ms_fused_multicast": t_mc, "multicast_bitwise_equal": same_mc,
                      "ms_nccl_gather": t_nccl, "ms_no_gather": t_local,
                      "tflops_fused": 2.0 * M * N * K / t_fused / 1e9, "tflops_nccl": 2.0 * M * N * K / t_nccl / 1e9}))
all_same = torch.tensor([int(same and same_mc)], device=dev); dist.all_reduce(all_same, op=dist.ReduceOp.MIN)
dist.destroy_process_group()
sys.exit(0 if int(all_same.item()) == 1 else 1)

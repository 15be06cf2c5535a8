#!/usr/bin/env python
"""torchrun target: row-sharded quantisation broadcast over NVLS + gated GEMM (ColumnShardedFp8Linear.forward_sharded_quant)
against the replicated path (forward_local) — bitwise, over several steps with changing data — and the step time of both.
usage: torchrun --nproc-per-node N tools/sharded_quant_check.py [M N K]"""
import json, os, sys
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
gw = torch.Generator(device=dev).manual_seed(77 + rank)
wq = ((torch.rand(hi - lo, K, generator=gw, device=dev) - 0.5) * 896).to(torch.float8_e4m3fn)
ws = (torch.rand(-((hi - lo) // -128), K // 128, generator=gw, device=dev) + 0.5) * 1e-3
layer = ColumnShardedFp8Linear(sgl_kernel.Fp8BlockLinear(wq, ws), N)
ok = layer.sharded_quant_supported(M) and layer.enable_sharded_quant(M)
TRANSPORTS = ["copy", "multicast"] if (ok and layer._sq_mc) else ["copy"]
res = {"world": world, "M": M, "N": N, "K": K, "multicast_available": bool(ok)}
same = True
if ok:
    # five steps with different batches (the same batch on every rank): both buffer sets, counters accumulating
    for step in range(5):
        g = torch.Generator(device=dev).manual_seed(1000 + step)
        x = torch.randn(M, K, generator=g, device=dev).to(torch.bfloat16)
        ref = layer.forward_local(x)
        tr = TRANSPORTS[step % len(TRANSPORTS)]
        out = layer.forward_sharded_quant(x, transport=tr)
        torch.cuda.synchronize()
        eq = bool(torch.equal(out, ref))
        res[f"step{step}_{tr}"] = eq
        same = same and eq
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev).view(torch.int64)

    def timeit(fn, n=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        ts = []
        for _ in range(n):
            flush.sum()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        t = torch.tensor([sum(ts) / n], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    t_rep = timeit(lambda: layer.forward_local(x))
    t_shq = timeit(lambda: layer.forward_sharded_quant(x))
    if "multicast" in TRANSPORTS:
        res["ms_sharded_quant_step_multicast"] = timeit(lambda: layer.forward_sharded_quant(x, transport="multicast"))
    xq, xs = layer.local.quantize_input(x)
    t_gemm = timeit(lambda: layer.local.forward_quantized(xq, xs))
    from sgl_kernel import _ops
    zero = torch.zeros(M // 256, dtype=torch.int32, device=dev)
    for whole in (False, True):      # the gated kernel alone, every counter already at its target: cost of the walk order + polls
        res["ms_gated_gemm_alone_" + ("panels" if whole else "slices")] = timeit(
            lambda: _ops._gemm_gated(xq, wq.t(), xs, ws.t(), torch.bfloat16, zero, 0, rank, world, whole_panels=whole))
    res.update({"bitwise_equal_to_replicated": same, "ms_replicated_quant_step": t_rep, "ms_sharded_quant_step": t_shq,
                "ms_gemm_alone": t_gemm, "tflops_replicated": 2.0 * M * N * K / t_rep / 1e9,
                "tflops_sharded_quant": 2.0 * M * N * K / t_shq / 1e9})
flag = torch.tensor([int(same)], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    res["bitwise_equal_on_all_ranks"] = bool(flag.item())
    print(json.dumps(res), flush=True)
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) == 1 else 1)

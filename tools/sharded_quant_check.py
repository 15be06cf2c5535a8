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


def _gated(*args, **kw):
    """the gated GEMM with a bounded wait (a few seconds): here a trap is better than a box that hangs"""
    kw.setdefault("wait_polls", 4_000_000)
    return _ops._gemm_gated(*args, **kw)


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
layer.gated_wait_polls = 16_000_000
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

    def step_ms(fn):
        flush.sum()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        return e0.elapsed_time(e1)

    from sgl_kernel import _ops
    xq, xs = layer.local.quantize_input(x)
    zero = torch.zeros(M // 256, dtype=torch.int32, device=dev)
    fns = {"replicated_quant_step": lambda: layer.forward_local(x),
           "sharded_quant_step": lambda: layer.forward_sharded_quant(x),
           "gemm_alone": lambda: layer.local.forward_quantized(xq, xs),
           "gated_gemm_alone": lambda: _gated(xq, wq.t(), xs, ws.t(), torch.bfloat16, zero, 0, rank, world, whole_panels=True)}
    if "multicast" in TRANSPORTS:
        fns["sharded_quant_step_multicast"] = lambda: layer.forward_sharded_quant(x, transport="multicast")
    # interleaved rounds (the clocks drift under the power cap); every configuration is a collective or rank-local, and
    # every rank runs them in the same order
    for f in fns.values():
        f(); f()
    torch.cuda.synchronize(); dist.barrier()
    ts = {k: [] for k in fns}
    for _ in range(15):
        for k, f in fns.items():
            ts[k].append(step_ms(f))
    for k, v in ts.items():
        t = torch.tensor([sorted(v)[len(v) // 2]], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res["ms_" + k] = round(float(t.item()), 4)
    res["bitwise_equal_to_replicated"] = same
    res["tflops_replicated"] = round(2.0 * M * N * K / res["ms_replicated_quant_step"] / 1e9, 1)
    res["tflops_sharded_quant"] = round(2.0 * M * N * K / res["ms_sharded_quant_step"] / 1e9, 1)
flag = torch.tensor([int(same)], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    res["bitwise_equal_on_all_ranks"] = bool(flag.item())
    print(json.dumps(res), flush=True)
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) == 1 else 1)

"""torchrun worker of tests/test_zy_multigpu_gpu.py: every multi-GPU form of the N-sharded layer against the plain
per-rank layer call (forward_local) and its NCCL gather — bitwise, over several steps, with the ranks deliberately out of
step with each other.  One process per GPU, NCCL for the plumbing.  No reference counterpart (the reference has no
communication layer).  Prints one JSON line on rank 0; exit code != 0 on any mismatch."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
import torch.distributed as dist

import sgl_kernel
from sgl_kernel.sharded import ColumnShardedFp8Linear, HostStreamedShardedFp8Linear, shard_bounds

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
M, N, K = 256 * world * 3, 2048 + 128, 2048          # ragged last panel when N % (128 * world) != 0
lo, hi = shard_bounds(N, world, rank)
gw = torch.Generator(device=dev).manual_seed(77)       # same full weight on every rank, each keeps its panel
wq_full = ((torch.rand(N, K, generator=gw, device=dev) - 0.5) * 896).to(torch.float8_e4m3fn)
ws_full = (torch.rand(-(N // -128), K // 128, generator=gw, device=dev) + 0.5) * 1e-3
full = sgl_kernel.Fp8BlockLinear(wq_full, ws_full)
layer = ColumnShardedFp8Linear(sgl_kernel.Fp8BlockLinear(wq_full[lo:hi].contiguous(), ws_full[lo // 128:-(hi // -128)].contiguous()), N)
res, bad = {"world": world, "M": M, "N": N, "K": K}, []


def batch(step):
    g = torch.Generator(device=dev).manual_seed(1000 + step)          # the same batch on every rank
    return torch.randn(M, K, generator=g, device=dev).to(torch.bfloat16)


def skew(step):
    """a different rank is late at every step (~2 ms of spinning on its stream before it joins)"""
    if step % world == rank:
        torch.cuda._sleep(4_000_000)


def note(name, ok):
    res[name] = bool(ok)
    if not ok:
        bad.append(name)


# 0. the sharded layer against the unsharded one: gather(forward_local) == full layer, bit for bit
x = batch(0)
ref_full = full(x)
note("nccl_gather_equals_unsharded_layer", torch.equal(layer.gather(layer.forward_local(x)), ref_full))

# 1. GEMM with the all-gather fused into its epilogue, unicast stores and (if the fabric offers it) NVLS multicast
layer.enable_fused_gather(M)
forms = [("unicast", False)] + ([("multicast", True)] if layer._hdl.multicast_ptr else [])
res["multicast_available"] = len(forms) == 2
for name, mc in forms:
    ok = True
    for step in range(4):
        x = batch(step)
        want = full(x)
        skew(step)
        got = layer.forward_fused_gather(x, multicast=mc)
        torch.cuda.synchronize()
        ok = ok and torch.equal(got, want)
        dist.barrier()                                               # nobody overwrites the buffer while a peer still compares
    note(f"fused_gather_{name}", ok)

# 2. row-sharded quantisation + gated GEMM, both transports, consecutive steps with no barrier between them
if layer.sharded_quant_supported(M) and layer.enable_sharded_quant(M):
    layer.gated_wait_polls = 16_000_000                              # a lost rank fails the test instead of hanging the box
    transports = ["copy"] + (["multicast"] if layer._sq_mc else [])
    for tr in transports:
        outs, wants = [], []
        for step in range(6):                                        # both buffer sets several times, counters accumulating
            x = batch(10 + step)
            wants.append(layer.forward_local(x))
            skew(step)
            outs.append(layer.forward_sharded_quant(x, transport=tr))
        torch.cuda.synchronize()
        note(f"sharded_quant_{tr}", all(torch.equal(o, w) for o, w in zip(outs, wants)))
        dist.barrier()
else:
    res["sharded_quant"] = "not supported for this shape"

# 3. host buffers in, host buffers out
runner = HostStreamedShardedFp8Linear(layer, M)
xs = [batch(20 + i).cpu().pin_memory() for i in range(3)]
ys = [torch.empty(M, hi - lo, dtype=torch.bfloat16).pin_memory() for _ in range(3)]
for i in range(3):
    skew(i)
    runner.submit(xs[i], ys[i])
runner.synchronize()
note("host_streamed_sharded", all(torch.equal(ys[i], layer.forward_local(xs[i].to(dev)).cpu()) for i in range(3)))

flag = torch.tensor([len(bad)], device=dev)
dist.all_reduce(flag)
if rank == 0:
    res["failed_on_some_rank"] = int(flag.item())
    print(json.dumps(res), flush=True)
if bad:
    print(f"rank {rank}: MISMATCH in {bad}", file=sys.stderr, flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(1 if int(flag.item()) else 0)

#!/usr/bin/env python
"""One GPU: the gated GEMM with every arrival counter already at its target against the plain GEMM on the same buffers —
what the walk order and the arrival polls of sglk_fp8_blockwise_scaled_mm_gated cost by themselves.
usage: gated_ab.py [worlds=2,4,8] [M=16384] [N_total=18432] [K=7168]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
import sgl_kernel
from sgl_kernel import _ops


def _gated(*args, **kw):
    """the gated GEMM with a bounded wait (a few seconds): here a trap is better than a box that hangs"""
    kw.setdefault("wait_polls", 4_000_000)
    return _ops._gemm_gated(*args, **kw)

dev = "cuda"
worlds = [int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "2,4,8").split(",")]
M, NT, K = (int(sys.argv[i]) if len(sys.argv) > i else d for i, d in ((2, 16384), (3, 18432), (4, 7168)))
x = torch.randn(M, K, device=dev).to(torch.bfloat16)
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev).view(torch.int64)


def burst_ms(fn, burst=6):
    flush.sum()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fn()
    e0.record()
    for _ in range(burst):
        fn()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / burst


for world in worlds:
    N = NT // world
    wq = ((torch.rand(N, K, device=dev) - 0.5) * 896).to(torch.float8_e4m3fn)
    ws = (torch.rand(-(N // -128), K // 128, device=dev) + 0.5) * 1e-3
    layer = sgl_kernel.Fp8BlockLinear(wq, ws)
    xq, xs = layer.quantize_input(x)
    zero = torch.zeros(M // 256, dtype=torch.int32, device=dev)
    ref = layer.forward_quantized(xq, xs)
    fns = {"plain": lambda: layer.forward_quantized(xq, xs)}
    for whole in (False, True):
        for rank in sorted({0, world - 1}):
            out = _gated(xq, wq.t(), xs, ws.t(), torch.bfloat16, zero, 0, rank, world, whole_panels=whole)
            assert torch.equal(out, ref)
            fns[f"gated_{'panels' if whole else 'slices'}_rank{rank}"] = (
                lambda rank=rank, whole=whole: _gated(xq, wq.t(), xs, ws.t(), torch.bfloat16, zero, 0, rank, world,
                                                                whole_panels=whole))
    # interleaved rounds: the chip's clock drifts under its power cap, so configurations are compared round by round
    ts = {k: [] for k in fns}
    for _ in range(2):
        for f in fns.values():
            f()
    for _ in range(9):
        for k, f in fns.items():
            ts[k].append(burst_ms(f))
    rec = {"world": world, "M": M, "N_local": N, "K": K}
    rec.update({"ms_" + k: round(sorted(v)[len(v) // 2], 4) for k, v in ts.items()})
    print(json.dumps(rec), flush=True)

#!/usr/bin/env python
"""GPU-side time of small launches: 20 launches captured in one CUDA graph, replayed (hides Python/launch overhead)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
import sgl_kernel
from sgl_kernel import _ops
dev = "cuda"

def graph_time(fn, n=20, reps=10):
    # warm up ON the capture stream: the stream-K workspace is cached per (device, stream) and is never created inside
    # a capture (sgl_kernel/_ops.py _get_workspace) — without this the captured launches run whole tiles only
    g = torch.cuda.CUDAGraph(); st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        fn()
    torch.cuda.synchronize()
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            for _ in range(n): fn()
    g.replay(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1) / n)
    ts.sort(); return ts[len(ts) // 2] * 1e3

def main():
    for m in (1, 16, 128, 512, 1024):
        x = torch.randn(m, 7168, dtype=torch.bfloat16, device=dev)
        q = torch.empty(m, 7168, dtype=torch.float8_e4m3fn, device=dev)
        s = torch.empty(m, 56, device=dev)
        us = graph_time(lambda: sgl_kernel.sgl_per_token_group_quant_8bit(x, q, s, 128, 1e-10, -448.0, 448.0, enable_v2=False))
        print(json.dumps({"op": "quant", "M": m, "us_graph": round(us, 2), "GBps": round(21728 * m / us / 1e3, 1)}))
    for N in (2112, 24576):
        for M in (1, 16, 64, 128, 256):
            K = 7168
            a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
            bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
            sa = (torch.randn(K // 128, M, device=dev) * 1e-3).t()
            sb = torch.randn(K // 128, -(N // -128), device=dev) * 1e-3
            for sk in (True, False):
                us = graph_time(lambda: _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, -1, 0, sk))
                print(json.dumps({"op": "gemm", "M": M, "N": N, "K": K, "stream_k": sk, "us_graph": round(us, 2),
                                  "GBps_weights": round(N * K / us / 1e3, 1), "note": "L2-warm for N=2112 (15 MB)"}))


if __name__ == '__main__':
    main()

#!/usr/bin/env python
"""Fixed cost / per-K-block cost of the decode GEMM: N = 148*128*t (t whole tiles per CTA, no ragged wave), whole-tile
vs forced stream-K schedules, weights alternating between two buffers so that no launch finds its weights in L2.
usage: decode_fit.py [variants=30,31] [Ms=16,128] [K=7168]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200"), os.path.join(ROOT, "tools")]
import torch
from sgl_kernel import _ops
dev = "cuda"
variants = [int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "30,31").split(",")]
Ms = [int(v) for v in (sys.argv[2] if len(sys.argv) > 2 else "16,128").split(",")]
K = int(sys.argv[3]) if len(sys.argv) > 3 else 7168


def graph_time2(fns, n=10, reps=8):
    for f in fns: f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph(); st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            for i in range(n): fns[i % len(fns)]()
    g.replay(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1) / n)
    ts.sort(); return ts[len(ts) // 2] * 1e3


for t in (1, 2, 4, 8):
    N = 148 * 128 * t
    bts = [(torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn) for _ in range(2 if N * K < (400 << 20) else 2)]
    sb = (torch.randn(-(N // -128), K // 128, device=dev) * 1e-3).t()
    for M in Ms:
        a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
        sa = torch.randn(M, K // 128, device=dev) * 1e-3
        for v in variants:
            for sk in (False, "force"):
                with _ops.stream_k_mode("force" if sk else "auto"):
                    fns = [(lambda b=b: _ops._gemm(a, b.t(), sa, sb, torch.bfloat16, v, 0, sk if sk else False)) for b in bts]
                    us = graph_time2(fns)
                print(json.dumps({"tiles_per_cta": t, "N": N, "K": K, "M": M, "variant": v, "stream_k": bool(sk), "us": round(us, 2),
                                  "us_per_kblock": round(us / (t * K / 128), 4), "GBps_w": round(N * K / us / 1e3, 1)}), flush=True)
    del bts

#!/usr/bin/env python
"""GPU-side time (20 launches per CUDA graph, L2 holds at most 126 MB of the 176 MB weight) of the decode GEMM for the
old 128-wide-tile kernel (variant 0) and the swap-AB kernel (30: single CTA, 31: CTA pair), DSv3 decode projections.
usage: decode_bench.py [variants=0,30,31] [Ms=1,16,32,64,128,256]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200"), os.path.join(ROOT, "tools")]
import torch
from sgl_kernel import _ops
from graph_time import graph_time
from oracle import torch_oracle as O
dev = "cuda"
variants = [int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "0,30,31").split(",")]
Ms = [int(v) for v in (sys.argv[2] if len(sys.argv) > 2 else "1,16,32,64,128,256").split(",")]
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
K = 7168
for N in (2112, 24576):
    bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
    sb = (torch.randn(-(N // -128), K // 128, device=dev) * 1e-3).t()
    for M in Ms:
        a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
        sa = torch.randn(M, K // 128, device=dev) * 1e-3
        ref = None
        for v in variants:
            try:
                out = _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, v, 0, True)
                torch.cuda.synchronize()
            except RuntimeError as e:
                print(json.dumps({"M": M, "N": N, "variant": v, "error": str(e)[:80]})); continue
            if ref is None:
                ref = out.float()
            err = float((out.float() - ref).abs().max() / ref.abs().max())
            us = sorted(graph_time(lambda: _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, v, 0, True)) for _ in range(3))[1]
            by = M * K + N * K + 2 * M * N + 4 * (M * (K // 128) + (K // 128) * -(N // -128))
            print(json.dumps({"M": M, "N": N, "variant": v, "us_graph": round(us, 2), "GBps": round(by / us / 1e3, 1),
                              "frac_hbm": round(by / us / 1e3 / peaks["hbm_gbs"], 3), "maxdiff_vs_first": round(err, 6)}), flush=True)

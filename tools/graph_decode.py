#!/usr/bin/env python
"""GPU-side time of the decode GEMM (20 launches in one CUDA graph, row-major scales, workspace on): the figure that
matters inside a CUDA-graph decode loop.  usage: graph_decode.py [tag]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
from sgl_kernel import _ops
from graph_time import graph_time
dev = "cuda"
tag = sys.argv[1] if len(sys.argv) > 1 else ""
K = 7168
for N in (2112, 24576):
    for M in (1, 16, 64, 128, 256):
        a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
        bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
        sa = torch.randn(M, K // 128, device=dev) * 1e-3
        sb = (torch.randn(-(N // -128), K // 128, device=dev) * 1e-3).t()
        us = sorted(graph_time(lambda: _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, -1, 0, True)) for _ in range(3))[1]
        print(json.dumps({"tag": tag, "M": M, "N": N, "K": K, "us_graph": round(us, 2), "GBps_weights": round(N * K / us / 1e3, 1)}))

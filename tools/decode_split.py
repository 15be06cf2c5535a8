#!/usr/bin/env python
"""N=2112 (fewer tiles than SMs): workers per tile of the decode kernel's stream-K split vs GPU-side time."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200"), os.path.join(ROOT, "tools")]
import torch
from sgl_kernel import _ops
from graph_time import graph_time
dev = "cuda"
K = 7168
for N in (2112, 4096, 7168):
    bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
    sb = (torch.randn(-(N // -128), K // 128, device=dev) * 1e-3).t()
    for M in (16, 64):
        a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
        sa = torch.randn(M, K // 128, device=dev) * 1e-3
        for v in (30, 31):
            row = {}
            for split in (1, 2, 3, 4, 6, 8, 12, 0):
                us = sorted(graph_time(lambda: _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, v, 0, True, 0, split)) for _ in range(3))[1]
                row[split] = round(us, 2)
            print(json.dumps({"N": N, "M": M, "variant": v, "us_by_workers_per_tile": row}), flush=True)

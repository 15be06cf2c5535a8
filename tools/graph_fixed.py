#!/usr/bin/env python
"""GPU-side fixed cost of the GEMM launch: tiny problems in a CUDA graph (no host overhead), K and N varied."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
from sgl_kernel import _ops
from graph_time import graph_time
dev = "cuda"
for (M, N, K, v) in [(64, 128, 128, 0), (64, 18944, 128, 0), (64, 18944, 512, 0), (64, 18944, 2048, 0), (64, 18944, 128, 1), (64, 18944, 2048, 1),
                     (64, 2112, 7168, 0), (256, 18944, 128, 3), (256, 18944, 2048, 3)]:
    a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
    bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
    sa = torch.randn(M, K // 128, device=dev) * 1e-3
    sb = (torch.randn(-(N // -128), K // 128, device=dev) * 1e-3).t()
    for sk in (False, True):
        us = graph_time(lambda: _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, v, 0, sk))
        print(json.dumps({"M": M, "N": N, "K": K, "variant": v, "workspace": sk, "us_graph": round(us, 2)}))

#!/usr/bin/env python
"""One launch per (variant, M) of the decode GEMM at N=24576 K=7168 for ncu.  usage: ncu_decode.py 30:128,31:16 [N]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
from sgl_kernel import _ops
dev = "cuda"
cases = [tuple(int(v) for v in c.split(":")) for c in sys.argv[1].split(",")]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 24576
K = 7168
bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
sb = (torch.randn(-(N // -128), K // 128, device=dev) * 1e-3).t()
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev).view(torch.int64)
for v, M in cases:
    a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
    sa = torch.randn(M, K // 128, device=dev) * 1e-3
    for _ in range(2):
        flush.sum()
        _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, v, 0, True)
    torch.cuda.synchronize()
print("ok")

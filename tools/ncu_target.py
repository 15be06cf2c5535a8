#!/usr/bin/env python
"""Tiny launcher for ncu captures: `ncu_target.py quant [M]`, `ncu_target.py tokenquant [M]` or `ncu_target.py gemm M N K [variant]`."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch  # noqa: E402

import sgl_kernel  # noqa: E402
from sgl_kernel import _ops  # noqa: E402

dev = "cuda"
what = sys.argv[1]
if what == "quant":
    m = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
    col = len(sys.argv) > 3 and sys.argv[3] == "col"
    x = torch.randn(m, 7168, dtype=torch.bfloat16, device=dev)
    q = torch.empty(m, 7168, dtype=torch.float8_e4m3fn, device=dev)
    s = sgl_kernel.create_per_token_group_quant_fp8_output_scale((m, 7168), dev, 128, col, col, False)
    for _ in range(3):
        sgl_kernel.sgl_per_token_group_quant_8bit(x, q, s, 128, 1e-10, -448.0, 448.0, enable_v2=False)
elif what == "tokenquant":
    m = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
    x = torch.randn(m, 7168, dtype=torch.bfloat16, device=dev)
    q = torch.empty(m, 7168, dtype=torch.float8_e4m3fn, device=dev)
    s = torch.zeros(m, device=dev)
    for _ in range(3):
        sgl_kernel.sgl_per_token_quant_fp8(x, q, s)
else:
    M, N, K = (int(v) for v in sys.argv[2:5])
    v = int(sys.argv[5]) if len(sys.argv) > 5 else -1
    a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
    bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
    sa = (torch.randn(K // 128, M, device=dev) * 1e-3).t()
    sb = torch.randn(K // 128, -(N // -128), device=dev) * 1e-3
    for _ in range(3):
        _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, v)
torch.cuda.synchronize()
print("done")

#!/usr/bin/env python
"""Bitwise comparison of two GEMM variants on a list of shapes (whole-tile and stream-K schedules).
usage: variant_eq.py vA vB [MxNxK ...]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
from sgl_kernel import _lib, _ops
from oracle import torch_oracle as O
dev = "cuda"
va, vb = int(sys.argv[1]), int(sys.argv[2])
shapes = sys.argv[3:] or ["256x256x512", "300x512x1024", "129x768x256", "1000x1280x7168", "4096x4096x2048", "384x2304x7168", "8192x8192x8192"]
lib = _lib.load()
for sh in shapes:
    M, N, K = (int(v) for v in sh.split("x"))
    a = ((torch.rand(M, K, device=dev) - 0.5) * 896).to(torch.float8_e4m3fn)
    bt = ((torch.rand(N, K, device=dev) - 0.5) * 896).to(torch.float8_e4m3fn)
    sa = torch.randn(M, K // 128, device=dev) * 1e-3
    sb = (torch.randn(-(N // -128), K // 128, device=dev) * 1e-3).t()
    for mode in ("never", "force"):
        with _ops.stream_k_mode(mode):
            ya = _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, va, 0, True)
            yb = _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, vb, 0, True)
            torch.cuda.synchronize()
        nbad = int((ya.view(torch.int16) != yb.view(torch.int16)).sum())
        ref_ok = ""
        if M * N * K <= 2e10:
            ref = O.baseline_scaled_mm(a.cpu(), bt.cpu().t(), sa.cpu(), sb.cpu(), torch.bfloat16)
            ref_ok = f" rel_err_vs_oracle={O.rel_err(yb.cpu(), ref):.2e}"
        print(f"{sh} stream_k_mode={mode}: {nbad} / {ya.numel()} elements differ{ref_ok}", flush=True)

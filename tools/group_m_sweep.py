#!/usr/bin/env python
"""Rasterisation sweep of the dense GEMM: m-blocks per raster group (the SGLK_GEMM_GROUP_M_SHIFT bits of the flags argument), interleaved round-robin
under the power cap, median per setting.  usage: group_m_sweep.py MxNxK g1,g2,... [rounds]   (ONCE=1: one launch per
setting, for `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum`)"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
from sgl_kernel import _lib, _ops
dev = "cuda"
M, N, K = (int(v) for v in sys.argv[1].split("x"))
groups = [int(v) for v in sys.argv[2].split(",")]
rounds = int(sys.argv[3]) if len(sys.argv) > 3 else 8
lib = _lib.load()
a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
sa = torch.randn(M, K // 128, device=dev) * 1e-3
sb = (torch.randn(-(N // -128), K // 128, device=dev) * 1e-3).t()
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev).view(torch.int64)
ref = None
for g in groups:
    flush.sum()
    out = _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, -1, 0, True, g)
    torch.cuda.synchronize()
    if ref is None:
        ref = out.clone()
    else:
        assert torch.equal(ref, out), f"group_m={g} changed the result"
if os.environ.get("ONCE") == "1":
    sys.exit(0)
times = {g: [] for g in groups}
for r in range(rounds + 2):
    for g in groups:
        for _ in range(3):
            flush.sum()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, -1, 0, True, g)
            e1.record(); e1.synchronize()
            if r >= 2:
                times[g].append(e0.elapsed_time(e1))
for g in groups:
    t = sorted(times[g]); med = t[len(t) // 2]
    print(json.dumps({"shape": sys.argv[1], "group_m": g, "us_med": round(med * 1e3, 1),
                      "TFLOPS_med": round(2.0 * M * N * K / med / 1e9, 1), "TFLOPS_best": round(2.0 * M * N * K / t[0] / 1e9, 1)}))

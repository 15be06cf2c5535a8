#!/usr/bin/env python
"""A/B timing of GEMM variants under the power cap: variants are interleaved round-robin (so they see the same
thermal state) and the per-variant median over all rounds is reported.  usage: ab_gemm.py MxNxK v1,v2,... [rounds]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
from sgl_kernel import _lib, _ops
dev = "cuda"
M, N, K = (int(v) for v in sys.argv[1].split("x"))
variants = [int(v) for v in sys.argv[2].split(",")]
rounds = int(sys.argv[3]) if len(sys.argv) > 3 else 8
a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
if os.environ.get("SCALES", "col") == "row":
    sa = torch.randn(M, K // 128, device=dev) * 1e-3
    sb = (torch.randn(-(N // -128), K // 128, device=dev) * 1e-3).t()
else:
    sa = (torch.randn(K // 128, M, device=dev) * 1e-3).t()
    sb = torch.randn(K // 128, -(N // -128), device=dev) * 1e-3
REPEAT = int(os.environ.get('REPEAT', '1'))   # launches inside one timed region (reported time is the total)
SK = os.environ.get("NOWS", "0") != "1"   # NOWS=1: no workspace (no stream-K, no scale re-layout)
NUM_SMS = int(os.environ.get('NUM_SMS', '0'))   # restrict the persistent grid (per-SM efficiency without L2 / power limits)
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev).view(torch.int64)
times = {v: [] for v in variants}
for v in variants:
    _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, v, NUM_SMS, SK)
torch.cuda.synchronize()
for r in range(rounds + 2):
    for v in variants:
        for _ in range(3):
            # NOFLUSH=1: keep L2 warm, but still put ~100 us of device work in front so the host-side launch cost stays hidden
            flush.sum() if os.environ.get('NOFLUSH', '0') != '1' else torch.cuda._sleep(200000)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(REPEAT): _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, v, NUM_SMS, SK)
            e1.record(); e1.synchronize()
            if r >= 2:
                times[v].append(e0.elapsed_time(e1))
lib = _lib.load()
for v in variants:
    t = sorted(times[v]); med = t[len(t) // 2]
    print(json.dumps({"shape": sys.argv[1], "variant": v, "name": lib.sglk_gemm_variant_name(v).decode() if v >= 0 else "auto",
                      "us_med": round(med * 1e3, 1), "TFLOPS_med": round(2.0 * M * N * K / med / 1e9, 1),
                      "TFLOPS_best": round(2.0 * M * N * K / t[0] / 1e9, 1), "TFLOPS_worst": round(2.0 * M * N * K / t[-1] / 1e9, 1)}))

#!/usr/bin/env python
"""Group-quant bandwidth at K=7168 (BASELINE config 2): CUDA events, read-only L2 flush between reps, median."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
import sgl_kernel
dev = "cuda"
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev).view(torch.int64)
for M in (4096, 16384, 32768):
    for col in (False, True):
        x = torch.randn(M, 7168, dtype=torch.bfloat16, device=dev)
        q = torch.empty(M, 7168, dtype=torch.float8_e4m3fn, device=dev)
        s = sgl_kernel.create_per_token_group_quant_fp8_output_scale((M, 7168), dev, 128, col, False, False)
        ts = []
        for i in range(25):
            flush.sum()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); sgl_kernel.sgl_per_token_group_quant_8bit(x, q, s, 128, 1e-10, -448.0, 448.0, enable_v2=False); e1.record()
            e1.synchronize()
            if i >= 5: ts.append(e0.elapsed_time(e1))
        ts.sort(); t = ts[len(ts) // 2]
        print(json.dumps({"M": M, "column_major_scales": col, "us": round(t * 1e3, 1), "GBps": round(21728 * M / t / 1e6, 1)}))

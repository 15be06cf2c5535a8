#!/usr/bin/env python
"""sgl_per_token_quant_fp8 bandwidth (the reference's bench grid, benchmark/bench_per_token_quant_fp8.py): CUDA events,
read-only L2 flush between reps, median.  Algorithmic bytes = tokens * (hidden * 3 + 4)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
import sgl_kernel
dev = "cuda"
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev).view(torch.int64)
for tokens, hidden in [(8192, 2048), (8192, 4096), (16384, 7168), (32768, 7168), (65536, 2048), (512, 7168), (128, 4096)]:
    x = torch.randn(tokens, hidden, dtype=torch.bfloat16, device=dev)
    q = torch.empty(tokens, hidden, dtype=torch.float8_e4m3fn, device=dev)
    s = torch.zeros(tokens, device=dev)
    ts = []
    for i in range(25):
        flush.sum()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); sgl_kernel.sgl_per_token_quant_fp8(x, q, s); e1.record()
        e1.synchronize()
        if i >= 5: ts.append(e0.elapsed_time(e1))
    ts.sort(); t = ts[len(ts) // 2]
    print(json.dumps({"tokens": tokens, "hidden": hidden, "us": round(t * 1e3, 1),
                      "GBps": round(tokens * (hidden * 3 + 4) / t / 1e6, 1)}))

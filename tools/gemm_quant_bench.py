#!/usr/bin/env python
"""Fused GEMM+quant epilogue vs the separate ops at the bench shape: CUDA events, read-only L2 flush, median."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
import sgl_kernel
dev = "cuda"
M, N, K = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "16384x18432x7168").split("x"))
a = ((torch.rand(M, K, device=dev) - 0.5) * 896).to(torch.float8_e4m3fn)
bt = ((torch.rand(N, K, device=dev) - 0.5) * 896).to(torch.float8_e4m3fn)
sa = torch.rand(M, K // 128, device=dev) * 1e-3 + 5e-4
sbt = torch.rand(N // 128, K // 128, device=dev) * 1e-2 + 5e-3
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev).view(torch.int64)

def sep_plain():
    y = sgl_kernel.fp8_blockwise_scaled_mm(a, bt.t(), sa, sbt.t(), torch.bfloat16)
    return sgl_kernel.per_token_group_quant_fp8(y, 128, enable_v2=False)
def sep_silu():
    y = sgl_kernel.fp8_blockwise_scaled_mm(a, bt.t(), sa, sbt.t(), torch.bfloat16)
    return sgl_kernel.per_token_group_quant_fp8(y, 128, fuse_silu_and_mul=True, enable_v2=True)
cases = {
    "gemm_only_bf16_out": lambda: sgl_kernel.fp8_blockwise_scaled_mm(a, bt.t(), sa, sbt.t(), torch.bfloat16),
    "separate: gemm + group quant": sep_plain,
    "fused: gemm_quant": lambda: sgl_kernel.fp8_blockwise_scaled_mm_quant(a, bt.t(), sa, sbt.t(), torch.bfloat16),
    "separate: gemm + silu_mul quant (v2 op)": sep_silu,
    "fused: gemm_quant silu_mul": lambda: sgl_kernel.fp8_blockwise_scaled_mm_quant(a, bt.t(), sa, sbt.t(), torch.bfloat16, fuse_silu_and_mul=True),
}
times = {k: [] for k in cases}
for r in range(10):
    for k, f in cases.items():
        flush.sum()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); e1.synchronize()
        if r >= 2: times[k].append(e0.elapsed_time(e1))
for k, t in times.items():
    t.sort(); med = t[len(t) // 2]
    print(json.dumps({"shape": f"{M}x{N}x{K}", "case": k, "ms_med": round(med, 4), "TFLOPS": round(2.0 * M * N * K / med / 1e9, 1)}))

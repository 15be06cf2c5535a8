#!/usr/bin/env python
"""Device-only timing sweep (CUDA events, L2 flushed between reps): quant GB/s over M, GEMM TFLOPS per
variant on the BASELINE shapes.  Writes one JSON object per line to stdout."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch  # noqa: E402

import sgl_kernel  # noqa: E402
from sgl_kernel import _lib, _ops  # noqa: E402


def timeit(fn, reps=20, warm=3, flush=None):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        if flush is not None:
            flush.sum()   # READ-only L2 flush: leaves clean lines (a memset would leave 126 MB of dirty lines to write back under the timed kernel)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="quant,gemm")
    ap.add_argument("--variants", default="")
    ap.add_argument("--shapes", default="")
    args = ap.parse_args()
    dev = "cuda"
    flush = torch.zeros(256 << 20, dtype=torch.int64 if False else torch.uint8, device=dev).view(torch.int64)
    if "quant" in args.what:
        for colmajor in (False, True):
            for m in [1, 16, 128, 512, 2048, 4096, 8192, 16384, 32768]:
                x = torch.randn(m, 7168, dtype=torch.bfloat16, device=dev)
                q = torch.empty(m, 7168, dtype=torch.float8_e4m3fn, device=dev)
                s = sgl_kernel.create_per_token_group_quant_fp8_output_scale((m, 7168), dev, 128, colmajor, colmajor, False)
                fn = lambda: sgl_kernel.sgl_per_token_group_quant_8bit(x, q, s, 128, 1e-10, -448.0, 448.0, enable_v2=False)
                med, best = timeit(fn, flush=flush)
                by = 21728 * m
                print(json.dumps({"op": "quant", "M": m, "K": 7168, "col_major": colmajor, "us_med": med * 1e3,
                                  "us_min": best * 1e3, "GBps_med": by / med / 1e6, "GBps_min": by / best / 1e6}), flush=True)
    if "gemm" in args.what:
        lib = _lib.load()
        nv = lib.sglk_gemm_num_variants()
        variants = [int(v) for v in args.variants.split(",")] if args.variants else list(range(nv))
        shapes = [(8192, 28672, 8192), (16384, 18432, 7168), (4096, 7168, 2048), (256, 24576, 7168), (64, 24576, 7168),
                  (1, 24576, 7168), (256, 2112, 7168), (1, 2112, 7168)]
        if args.shapes:
            shapes = [tuple(int(v) for v in s.split("x")) for s in args.shapes.split(",")]
        for (M, N, K) in shapes:
            a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
            bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
            sa = (torch.randn(K // 128, M, device=dev) * 1e-3).t()
            sb = torch.randn(K // 128, -(N // -128), device=dev) * 1e-3
            for v in variants + [-1]:
                try:
                    fn = lambda: _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, v)
                    med, best = timeit(fn, reps=10, flush=flush if M * N < (1 << 26) else None)
                except Exception as e:  # noqa: BLE001
                    print(json.dumps({"op": "gemm", "M": M, "N": N, "K": K, "variant": v, "error": str(e)[:200]}), flush=True)
                    continue
                fl = 2.0 * M * N * K
                by = M * K + N * K + 2 * M * N
                print(json.dumps({"op": "gemm", "M": M, "N": N, "K": K, "variant": v,
                                  "name": lib.sglk_gemm_variant_name(v).decode() if v >= 0 else "auto:" + lib.sglk_gemm_variant_name(lib.sglk_gemm_pick_variant(M, N, K)).decode(),
                                  "us_med": med * 1e3, "us_min": best * 1e3, "TFLOPS_med": fl / med / 1e9,
                                  "TFLOPS_min": fl / best / 1e9, "GBps_med": by / med / 1e6}), flush=True)
        # practical fp8 ceiling on this box: cuBLASLt per-tensor fp8 (library call, for the denominator only)
        try:
            n = 8192
            a = torch.randn(n, n, device=dev).to(torch.float8_e4m3fn)
            b = torch.randn(n, n, device=dev).to(torch.float8_e4m3fn).t()
            one = torch.ones((), device=dev)
            fn = lambda: torch._scaled_mm(a, b, scale_a=one, scale_b=one, out_dtype=torch.bfloat16)
            med, best = timeit(fn, reps=10)
            print(json.dumps({"op": "cublaslt_fp8_8192", "TFLOPS_med": 2.0 * n ** 3 / med / 1e9, "TFLOPS_min": 2.0 * n ** 3 / best / 1e9}), flush=True)
            a16, b16 = torch.randn(n, n, device=dev, dtype=torch.bfloat16), torch.randn(n, n, device=dev, dtype=torch.bfloat16)
            med, best = timeit(lambda: a16 @ b16, reps=10)
            print(json.dumps({"op": "cublas_bf16_8192", "TFLOPS_med": 2.0 * n ** 3 / med / 1e9, "TFLOPS_min": 2.0 * n ** 3 / best / 1e9}), flush=True)
            src = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
            dst = torch.empty_like(src)
            med, best = timeit(lambda: dst.copy_(src), reps=10)
            print(json.dumps({"op": "copy_1GiB", "GBps_med": 2 * (1 << 30) / med / 1e6, "GBps_min": 2 * (1 << 30) / best / 1e6}), flush=True)
        except Exception as e:  # noqa: BLE001
            print(json.dumps({"op": "ceilings", "error": str(e)[:300]}), flush=True)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Measure every BASELINE.json config on one B200 and print a markdown table + JSON lines.
Device timings: CUDA events, median of reps, read-only 256 MiB L2 flush between reps; parity verdict per
config from the CPU oracle on a row sample.  The CPU reference (config 1) is timed in the same run."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch  # noqa: E402

import sgl_kernel  # noqa: E402
from oracle import torch_oracle as O  # noqa: E402
from sgl_kernel import _lib, _ops  # noqa: E402

dev = "cuda"
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
HBM, FP8 = peaks["hbm_gbs"], 2 * peaks["bf16_tflops"]
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev).view(torch.int64)
rows = []


def timeit(fn, reps=15, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        flush.sum()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def emit(**kw):
    rows.append(kw)
    print(json.dumps(kw), flush=True)


# ---- config 1: CPU reference, M=128 N=4096 K=7168 ----
x, wq_t, ws_t = O.make_linear_inputs(128, 4096, 7168, seed=42)
O.fp8_linear_ref(x, wq_t, ws_t)
best = 1e9
for _ in range(5):
    t0 = time.perf_counter(); ref = O.fp8_linear_ref(x, wq_t, ws_t); best = min(best, time.perf_counter() - t0)
layer = sgl_kernel.Fp8BlockLinear(wq_t.t().contiguous().to(dev), ws_t.t().contiguous().to(dev))
xd = x.to(dev)
out = layer(xd); torch.cuda.synchronize()
ms = timeit(lambda: layer(xd))
emit(config="1 CPU reference chain M=128 N=4096 K=7168", cpu_ms=best * 1e3, cpu_threads=torch.get_num_threads(),
     cpu_gflops=2 * 128 * 4096 * 7168 / best / 1e9, gpu_us=ms * 1e3, gpu_tflops=2 * 128 * 4096 * 7168 / ms / 1e9,
     parity=f"rel_err {O.rel_err(out.cpu(), ref):.1e}")

# ---- config 2: quant sweep ----
for colmajor in (False, True):
    for m in [1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768]:
        torch.manual_seed(0)
        xc = torch.randn(m, 7168, dtype=torch.bfloat16)
        xg = xc.to(dev)
        q = torch.empty(m, 7168, dtype=torch.float8_e4m3fn, device=dev)
        s = sgl_kernel.create_per_token_group_quant_fp8_output_scale((m, 7168), dev, 128, colmajor, colmajor, False)
        fn = lambda: sgl_kernel.sgl_per_token_group_quant_8bit(xg, q, s, 128, 1e-10, -448.0, 448.0, enable_v2=False)
        ms = timeit(fn)
        idx = torch.arange(0, m, max(1, m // 64))
        qr, sr = O.per_token_group_quant_fp8_ref(xc[idx].contiguous(), 128)
        ok = torch.equal(q[idx.to(dev)].cpu().view(torch.uint8), qr.view(torch.uint8)) and torch.equal(s[idx.to(dev)].cpu(), sr)
        by = 21728 * m
        emit(config=f"2 quant K=7168 M={m} {'col' if colmajor else 'row'}-major scales", us=ms * 1e3, GBps=by / ms / 1e6,
             frac_measured_hbm=by / ms / 1e6 / HBM, frac_nominal_8TBs=by / ms / 1e6 / 8000, parity="bit-exact" if ok else "MISMATCH")


def gemm_case(tag, M, N, K):
    g = torch.Generator().manual_seed(1)
    a = ((torch.rand(M, K, generator=g) - 0.5) * 896).to(torch.float8_e4m3fn)
    bt = ((torch.rand(N, K, generator=g) - 0.5) * 896).to(torch.float8_e4m3fn)
    sa = torch.randn(M, K // 128, generator=g) * 1e-3
    sb = torch.randn(K // 128, -(N // -128), generator=g) * 1e-3
    ad, bd = a.to(dev), bt.to(dev).t()
    sad, sbd = sa.t().contiguous().to(dev).t(), sb.to(dev)
    out = sgl_kernel.fp8_blockwise_scaled_mm(ad, bd, sad, sbd, torch.bfloat16)
    torch.cuda.synchronize()
    idx = torch.unique(torch.cat([torch.arange(0, M, max(1, M // 48)), torch.tensor([M - 1])]))
    ref = O.baseline_scaled_mm(a[idx], bt.t(), sa[idx], sb, torch.bfloat16)
    r, c = O.rel_err(out[idx.to(dev)].cpu(), ref), O.cosine(out[idx.to(dev)].cpu(), ref)
    ms = timeit(lambda: sgl_kernel.fp8_blockwise_scaled_mm(ad, bd, sad, sbd, torch.bfloat16), reps=12)
    fl = 2.0 * M * N * K
    by = M * K + N * K + 2 * M * N + 4 * (M * (K // 128) + (K // 128) * -(N // -128))
    emit(config=f"{tag} M={M} N={N} K={K}", us=ms * 1e3, TFLOPS=fl / ms / 1e9, GBps=by / ms / 1e6,
         frac_fp8_peak_2x_measured_bf16=fl / ms / 1e9 / FP8, frac_nominal_4500=fl / ms / 1e9 / 4500,
         frac_measured_hbm=by / ms / 1e6 / HBM, variant=_lib.load().sglk_gemm_variant_name(_lib.load().sglk_gemm_pick_variant(M, N, K)).decode(),
         parity=f"rel_err {r:.1e} cos {c:.6f}")


for N in (2112, 24576):
    for M in (1, 2, 4, 8, 16, 32, 64, 128, 256):
        gemm_case("3 DSv3 decode", M, N, 7168)
gemm_case("4 Llama-3-70B gate/up prefill", 8192, 28672, 8192)
gemm_case("5 DSv3 dense MLP prefill (1 GPU)", 16384, 18432, 7168)

print("\n| config | time | throughput | fraction of roofline | parity |")
print("|---|---|---|---|---|")
for r in rows:
    if "cpu_ms" in r:
        print(f"| {r['config']} | CPU {r['cpu_ms']:.1f} ms ({r['cpu_threads']} threads) / GPU {r['gpu_us']:.1f} us | CPU {r['cpu_gflops']:.0f} GFLOP/s / GPU {r['gpu_tflops']:.1f} TFLOP/s | — | {r['parity']} |")
    elif "TFLOPS" in r:
        print(f"| {r['config']} | {r['us']:.1f} us | {r['TFLOPS']:.1f} TFLOP/s, {r['GBps']:.0f} GB/s | tensor {r['frac_fp8_peak_2x_measured_bf16']:.2f} of 2x measured bf16 ({r['frac_nominal_4500']:.2f} of 4.5 PF); HBM {r['frac_measured_hbm']:.2f} of measured | {r['parity']} |")
    else:
        print(f"| {r['config']} | {r['us']:.1f} us | {r['GBps']:.0f} GB/s | {r['frac_measured_hbm']:.2f} of measured copy ({r['frac_nominal_8TBs']:.2f} of 8 TB/s) | {r['parity']} |")

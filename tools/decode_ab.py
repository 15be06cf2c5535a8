#!/usr/bin/env python
"""A/B of the decode GEMM variants with every launch streaming its weight from HBM: the launches of one CUDA graph
rotate over enough weight copies to exceed 2 x L2 (bench.py's config-3 method), GPU-side time per launch.
usage: decode_ab.py [variants=-1,0,30,31,32,32d] [Ms=1,16,32,64,128] [Ns=2112,24576] [Ks=7168]
("32d" = variant 32 with one 2-D TMA load per K block instead of one 3-D load per stage; "32s4" / "32s2" / "32s1" =
variant 32 with K split over clusters of 4 / 2 / not at all)"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
from sgl_kernel import _lib, _ops
dev = "cuda"
arg = lambda i, d: [int(v) for v in (sys.argv[i] if len(sys.argv) > i else d).split(",")]
variants = (sys.argv[1] if len(sys.argv) > 1 else "-1,0,30,31,32,32d").split(",")
Ms, Ns, Ks = arg(2, "1,16,32,64,128"), arg(3, "2112,24576"), arg(4, "7168")
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}


def run(a, bt, sa, sb, v):
    if v.endswith("d"):
        return _ops._gemm(a, bt, sa, sb, torch.bfloat16, int(v[:-1]), extra_flags=_lib.DECODE_2D_LOADS)
    if "s" in v:                                     # "32s4": variant 32 with K split over clusters of 4 (1 = no split)
        vid, ks = v.split("s")
        return _ops._gemm(a, bt, sa, sb, torch.bfloat16, int(vid), split=int(ks))
    return _ops._gemm(a, bt, sa, sb, torch.bfloat16, int(v))


def graph_us(fns, reps=9):
    g, st = torch.cuda.CUDAGraph(), torch.cuda.Stream()
    with torch.cuda.stream(st):
        for f in fns[:2]:
            f()
    torch.cuda.synchronize()
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            for f in fns:
                f()
    g.replay(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1) / len(fns))
    return sorted(ts)[len(ts) // 2] * 1e3


for K, N in [(K, N) for K in Ks for N in Ns]:
    copies = min(-(-(256 << 20) // (N * K)), 64)
    launches = copies * -(-16 // copies)
    bts = [(torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn) for _ in range(copies)]
    sb = (torch.randn(-(N // -128), K // 128, device=dev) * 1e-3).t()
    for M in Ms:
        a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
        sa = torch.randn(M, K // 128, device=dev) * 1e-3
        ref = _ops._gemm(a, bts[0].t(), sa, sb, torch.bfloat16, 0).float()
        by = M * K + N * K + 2 * M * N + 4 * (M * (K // 128) + (K // 128) * -(N // -128))
        for v in variants:
            try:
                out = run(a, bts[0].t(), sa, sb, v)
                torch.cuda.synchronize()
            except RuntimeError as e:
                print(json.dumps({"M": M, "N": N, "variant": v, "error": str(e)[:80]})); continue
            rel = float((out.float() - ref).norm() / ref.norm())
            us = graph_us([(lambda i=i: run(a, bts[i % copies].t(), sa, sb, v)) for i in range(launches)])
            print(json.dumps({"M": M, "N": N, "K": K, "variant": v, "us": round(us, 2), "GBps": round(by / us / 1e3, 1),
                              "frac_hbm": round(by / us / 1e3 / peaks["hbm_gbs"], 3), "rel_vs_tile_kernel": round(rel, 7)}), flush=True)
    del bts

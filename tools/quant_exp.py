#!/usr/bin/env python
"""A/B of the group-quant loop-structure variants (csrc/quant_exp.cu) against the product kernel: bit-equality of
codes and scales, then CUDA-event timing with a read-only L2 flush between reps (median)."""
import ctypes, os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
# laboratory tool: needs the diagnostics build (make -C sgl-kernel-xpu_b200/csrc diag)
os.environ.setdefault("SGLK_LIBRARY", os.path.join(ROOT, "sgl-kernel-xpu_b200", "lib", "libsglk_diag.so"))
import sgl_kernel
from sgl_kernel import _lib
lib = _lib.load()
fn = lib.sglk_debug_quant_exp
fn.restype = ctypes.c_int
fn.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int64, ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
dev = "cuda"
flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev).view(torch.int64)
variants = [int(v) for v in sys.argv[1].split(",")] if len(sys.argv) > 1 else list(range(1, 13))
Ms = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [4096, 16384, 32768]

def timed(f):
    ts = []
    for i in range(25):
        flush.sum()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); e1.synchronize()
        if i >= 5: ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]

for M in Ms:
    x = torch.randn(M, 7168, dtype=torch.bfloat16, device=dev)
    x[0, 3] = -0.0
    q0 = torch.empty(M, 7168, dtype=torch.float8_e4m3fn, device=dev)
    s0 = torch.empty(M, 56, dtype=torch.float32, device=dev)
    prod = lambda: sgl_kernel.sgl_per_token_group_quant_8bit(x, q0, s0, 128, 1e-10, -448.0, 448.0, enable_v2=False)
    prod()
    t = timed(prod)
    print(json.dumps({"M": M, "variant": "product", "us": round(t * 1e3, 1), "GBps": round(21728 * M / t / 1e6, 1)}), flush=True)
    st = torch.cuda.current_stream().cuda_stream
    for v in variants:
        for cps in ((16, 4, 0) if v in (1, 3, 7, 8, 11) else (0,)):
            q = torch.zeros_like(q0); s = torch.zeros_like(s0)
            run = lambda: fn(x.data_ptr(), q.data_ptr(), s.data_ptr(), M * 56, 1e-10, v, cps if cps else 99, st)
            rc = run(); torch.cuda.synchronize()
            assert rc == 0, rc
            same = torch.equal(q.view(torch.uint8), q0.view(torch.uint8)) and torch.equal(s.view(torch.int32), s0.view(torch.int32))
            t = timed(run)
            print(json.dumps({"M": M, "variant": v, "ctas_per_sm_cap": cps if cps else 99, "bit_equal": same,
                              "us": round(t * 1e3, 1), "GBps": round(21728 * M / t / 1e6, 1)}), flush=True)

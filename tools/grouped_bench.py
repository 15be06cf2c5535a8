#!/usr/bin/env python
"""fp8_blockwise_scaled_grouped_mm (MoE) throughput: E experts, tokens spread unevenly (multiples of 16 as in the
reference test, tests/test_fp8_blockwise_moe.py:66-171), DeepSeek-V3 expert shapes — gate|up N=4096 K=7168 and down
N=7168 K=2048.  CUDA events, L2 flushed between launches, median of 9; TFLOP/s over the rows that exist, and against the
dense op on one [sum_m, N, K] problem (same flops, one weight).
usage: grouped_bench.py [Es=8,32,64] [avg tokens per expert=512]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
import sgl_kernel
dev = "cuda"
# defaults; bench.py imports this module and sets them, the command line is read only under __main__ (an importer's argv
# is not ours: parsing it at import time turned bench.py's f4 sub-record into an error string in BENCH_r02.json)
Es = [8, 32, 64]
avg = 512
_flush = None


def med(fn, reps=9):
    global _flush
    if _flush is None:
        _flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev).view(torch.int64)
    flush = _flush
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        flush.sum()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


def records():
    out = []
    g = torch.Generator().manual_seed(0)
    for E in Es:
        # uneven routing: 16 * U[1, 2*avg/16) tokens per expert
        ms = (16 * torch.randint(1, max(2, 2 * avg // 16), (E,), generator=g)).tolist()
        sum_m = sum(ms)
        offs = torch.tensor([sum(ms[:i]) for i in range(E)], dtype=torch.int32, device=dev)
        for name, N, K in (("gate_up", 4096, 7168), ("down", 7168, 2048)):
            a = ((torch.rand(sum_m, K, device=dev) - 0.5) * 896).to(torch.float8_e4m3fn)
            bt = ((torch.rand(E, N, K, device=dev) - 0.5) * 896).to(torch.float8_e4m3fn)
            sa = torch.rand(sum_m, K // 128, device=dev) * 1e-3 + 5e-4
            sbt = torch.rand(E, N // 128, K // 128, device=dev) * 1e-3 + 5e-4
            sizes = torch.tensor([[m, N, K] for m in ms], dtype=torch.int32, device=dev)
            o = torch.empty(sum_m, N, dtype=torch.bfloat16, device=dev)
            i64 = lambda: torch.empty(E, dtype=torch.int64, device=dev)
            lay = lambda: torch.zeros(E, 5, dtype=torch.int32, device=dev)
            st = torch.full((E,), K, dtype=torch.int64, device=dev)
            cs = torch.full((E,), N, dtype=torch.int64, device=dev)
            ws = torch.empty(1 << 20, dtype=torch.uint8, device=dev)
            args = (o, i64(), i64(), i64(), i64(), i64(), a, bt.transpose(1, 2), sa, sbt.transpose(1, 2), st, st, cs, lay(), lay(),
                    sizes, offs, ws)
            t = med(lambda: sgl_kernel.fp8_blockwise_scaled_grouped_mm(*args))
            # parity spot check: expert 0 against the dense op
            ref0 = sgl_kernel.fp8_blockwise_scaled_mm(a[:ms[0]], bt[0].t(), sa[:ms[0]], sbt[0].t(), torch.bfloat16)
            same = float((o[:ms[0]].float() - ref0.float()).norm() / ref0.float().norm()) <= 1e-3   # (summation orders differ)
            td = med(lambda: sgl_kernel.fp8_blockwise_scaled_mm(a, bt[0].t(), sa, sbt[0].t(), torch.bfloat16))
            fl = 2.0 * sum_m * N * K
            by = sum_m * K + E * N * K + 2 * sum_m * N
            out.append({"experts": E, "proj": name, "sum_m": sum_m, "N": N, "K": K, "ms": round(t, 4),
                        "tflops": round(fl / t / 1e9, 1), "GBps": round(by / t / 1e6, 1),
                        "dense_same_flops_ms": round(td, 4), "dense_tflops": round(fl / td / 1e9, 1),
                        "expert0_matches_dense_op": same})
            del a, bt, o
    return out


if __name__ == "__main__":
    if len(sys.argv) > 1:
        Es = [int(v) for v in sys.argv[1].split(",")]
    if len(sys.argv) > 2:
        avg = int(sys.argv[2])
    for r in records():
        print(json.dumps(r), flush=True)

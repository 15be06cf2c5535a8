#!/usr/bin/env python
"""Run ONE GEMM variant on a few small problems against the CPU oracle (used under `timeout` so a
deadlocked kernel cannot take the whole GPU session down).  usage: variant_check.py <variant> [quick]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch  # noqa: E402

from oracle import torch_oracle as O  # noqa: E402
from sgl_kernel import _lib, _ops  # noqa: E402

v = int(sys.argv[1])
name = _lib.load().sglk_gemm_variant_name(v).decode()
shapes = [(128, 128, 128), (128, 256, 512), (256, 256, 1024), (5, 512, 1024), (300, 640, 2048), (512, 2112, 1024)]
worst = 0.0
for (M, N, K) in shapes:
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M + N + K)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    t0 = time.time()
    out = _ops._gemm(a.cuda(), b.t().contiguous().cuda().t(), sa.t().contiguous().cuda().t(), sb.cuda(), torch.bfloat16, v)
    torch.cuda.synchronize()
    out = out.cpu()
    r, c = O.rel_err(out, ref), O.cosine(out, ref)
    worst = max(worst, r)
    print(f"variant {v} {name:24s} {M:5d}x{N:5d}x{K:5d} rel_err={r:.3e} cos={c:.6f} finite={bool(torch.isfinite(out.float()).all())} ({time.time()-t0:.2f}s)", flush=True)
    if not (r <= 1e-2):
        # localise: which 32x32 blocks are wrong
        d = (out.float() - ref.float()).abs()
        bad = d > (0.02 * ref.float().abs() + 1.0)
        rows = bad.any(1).nonzero().flatten()
        cols = bad.any(0).nonzero().flatten()
        print(f"   bad rows {rows[:8].tolist()}..{rows[-4:].tolist()} ({len(rows)})  bad cols {cols[:8].tolist()}..{cols[-4:].tolist()} ({len(cols)})")
        print("   out[0,:8]", out[0, :8].float().tolist(), "\n   ref[0,:8]", ref[0, :8].float().tolist())
print("VARIANT", v, name, "OK" if worst <= 1e-2 else "FAIL", flush=True)
sys.exit(0 if worst <= 1e-2 else 1)

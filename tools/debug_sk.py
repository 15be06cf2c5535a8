import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
from oracle import torch_oracle as O
from sgl_kernel import _lib, _ops
M, N, K = (int(v) for v in sys.argv[1:4]); variant = int(sys.argv[4])
a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M + N + K)
ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16).float()
ad, bd = a.cuda(), b.t().contiguous().cuda().t()
sad, sbd = sa.t().contiguous().cuda().t(), sb.cuda()
for sk in (False, True, True):
    for sms in ([0] if len(sys.argv) < 6 else [int(v) for v in sys.argv[5].split(",")]):
        out = _ops._gemm(ad, bd, sad, sbd, torch.bfloat16, variant, sms, sk)
        torch.cuda.synchronize()
        o = out.cpu().float()
        bad = (o - ref).abs() > (0.02 * ref.abs() + 1.0)
        print(f"sk={sk} sms={sms} rel={O.rel_err(o, ref):.3e} bad={int(bad.sum())}")
        if bad.any():
            rows = bad.any(1).nonzero().flatten().tolist(); cols = bad.any(0).nonzero().flatten().tolist()
            def runs(x):
                r=[]; s=x[0]; p=x[0]
                for v in x[1:]:
                    if v!=p+1: r.append((s,p)); s=v
                    p=v
                r.append((s,p)); return r
            print("  bad row runs", runs(rows)[:12]); print("  bad col runs", runs(cols)[:12])
            r0,c0=rows[0],cols[0]
            print("  out/ref ratio sample", (o[r0, c0:c0+4]/ref[r0, c0:c0+4]).tolist())

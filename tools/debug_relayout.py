#!/usr/bin/env python
"""Where do the re-laid-out-scales result and the K-contiguous-scales result differ?  usage: debug_relayout.py M N K variant"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
from sgl_kernel import _ops
from oracle import torch_oracle as O
M, N, K, variant = (int(v) for v in sys.argv[1:5])
DEV = "cuda"
a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M + N)
ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
sa_mmaj = sa.t().contiguous().to(DEV).t()
sb_rows = sb.contiguous().to(DEV)
sa_row, sb_k = sa.contiguous().to(DEV), sb.t().contiguous().to(DEV).t()
def run(tag, *args, **kw):
    o = _ops._gemm(*args, **kw); torch.cuda.synchronize(); return o
outs = {}
for rep in range(3):
    outs[f"relayout{rep}"] = run("r", ad, bd, sa_mmaj, sb_rows, torch.bfloat16, variant)
    outs[f"vec{rep}"] = run("v", ad, bd, sa_row, sb_k, torch.bfloat16, variant)
    outs[f"nows{rep}"] = run("n", ad, bd, sa_mmaj, sb_rows, torch.bfloat16, variant, stream_k=False)
    outs[f"vec_nows{rep}"] = run("n", ad, bd, sa_row, sb_k, torch.bfloat16, variant, stream_k=False)
base = outs["vec0"]
for k, v in outs.items():
    d = (v.float() - base.float()).abs()
    nz = d.nonzero()
    msg = f"{k:12s} ndiff={nz.shape[0]:8d} maxabs={d.max().item():.4g}"
    if nz.shape[0]:
        rows, cols = nz[:, 0], nz[:, 1]
        msg += f" rows[{rows.min().item()},{rows.max().item()}] cols[{cols.min().item()},{cols.max().item()}] first={nz[0].tolist()}"
    print(msg)

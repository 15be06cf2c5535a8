import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
from sgl_kernel import _ops
from graph_time import graph_time  # noqa
dev = "cuda"
for (M, N) in ((1, 2112), (64, 2112)):
    K = 7168
    a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
    bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
    sa = (torch.randn(K // 128, M, device=dev) * 1e-3).t()
    sb = torch.randn(K // 128, -(N // -128), device=dev) * 1e-3
    for v in (0, 6):
        for sms in (17, 34, 51, 68, 102, 148):
            us = graph_time(lambda: _ops._gemm(a, bt.t(), sa, sb, torch.bfloat16, v, sms, True))
            print(json.dumps({"M": M, "N": N, "variant": v, "num_sms": sms, "us_graph": round(us, 2)}))

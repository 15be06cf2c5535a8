import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
from sgl_kernel import _lib, _ops
dev = "cuda"
M, N, K = 8192, 8192, 8192
a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
bt = (torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
ROWMAJOR = len(sys.argv) > 2 and sys.argv[2] == "row"
if ROWMAJOR:   # scales contiguous along K blocks (quant default layout / weight-scale view)
    sa = torch.randn(M, K // 128, device=dev) * 1e-3
    sb = (torch.randn(N // 128, K // 128, device=dev) * 1e-3).t()
else:          # the reference test's layout: sa M-major, sb [G, NB] row-major
    sa = (torch.randn(K // 128, M, device=dev) * 1e-3).t()
    sb = torch.randn(K // 128, N // 128, device=dev) * 1e-3
lib = _lib.load()
ws = torch.zeros(8 << 20, dtype=torch.uint8, device=dev)
out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
VAR = int(sys.argv[1]) if len(sys.argv) > 1 else 15
for it in range(3):
    rc = lib.sglk_fp8_blockwise_scaled_mm_ex(a.data_ptr(), bt.data_ptr(), sa.data_ptr(), sb.data_ptr(), out.data_ptr(), M, N, K, K, K, N,
                                             sa.stride(0), sa.stride(1), sb.stride(0), sb.stride(1), 0, VAR, 0, ws.data_ptr(), ws.numel(),
                                             torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rc
torch.cuda.synchronize()
d = ws[1 << 20:(1 << 20) + 64 * 8 * 8].view(torch.int64).view(64, 8).cpu()
t0 = int(d[8, 0])
print("kb: per-phase clocks of epilogue warp 0 / MMA thread of CTA 0")
for kb in range(8, 32):
    r = [int(v) - t0 for v in d[kb, :8]]
    nxt = [int(v) - t0 for v in d[kb + 1, :8]]
    print(f"{kb:3d}: mma[tempty_ok {r[0]:6d}]  epi[first_chunk {r[3]:6d} | 3 chunks {r[4]-r[3]:4d} | release {r[5]-r[4]:4d} | trywait+ld {r[7]-r[5]:4d} | last fma {r[6]-r[7]:4d} | cpwait {nxt[2]-r[6]:4d} | lds+prefetch {nxt[1]-nxt[2]:4d} | ld wait {nxt[3]-nxt[1]:4d}]  period {nxt[3]-r[3]:4d}")

import statistics
per_epi = [int(d[kb + 1, 3]) - int(d[kb, 3]) for kb in range(8, 60)]
per_mma = [int(d[kb + 1, 0]) - int(d[kb, 0]) for kb in range(8, 60)]
print(json.dumps({"variant": VAR, "median_clk_per_kblock_epilogue": statistics.median(per_epi), "median_clk_per_kblock_mma": statistics.median(per_mma),
                  "mean_clk_per_kblock": sum(per_mma) / len(per_mma), "mma_exec_clk": 512}))

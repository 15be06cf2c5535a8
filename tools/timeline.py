import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]
import torch
# laboratory tool: needs the diagnostics build (make -C sgl-kernel-xpu_b200/csrc diag)
os.environ.setdefault("SGLK_LIBRARY", os.path.join(ROOT, "sgl-kernel-xpu_b200", "lib", "libsglk_diag.so"))
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
NSM = int(sys.argv[3]) if len(sys.argv) > 3 else 0      # run on this many SMs only (0: all)
for it in range(3):
    rc = lib.sglk_fp8_blockwise_scaled_mm_ex(a.data_ptr(), bt.data_ptr(), sa.data_ptr(), sb.data_ptr(), out.data_ptr(), M, N, K, K, K, N,
                                             sa.stride(0), sa.stride(1), sb.stride(0), sb.stride(1), 0, VAR, NSM, ws.data_ptr(), ws.numel(),
                                             torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rc
torch.cuda.synchronize()
# wall-clock of one more launch -> the SM clock the kernel actually ran at (clock64 ticks / elapsed time)
import threading, time
try:
    import pynvml
    pynvml.nvmlInit(); _h = pynvml.nvmlDeviceGetHandleByIndex(0)
except Exception:
    _h = None
samples = []
def _sample():
    while not _stop:
        if _h is not None:
            samples.append((pynvml.nvmlDeviceGetClockInfo(_h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(_h) / 1000.0))
        time.sleep(0.002)
_stop = False
th = threading.Thread(target=_sample); th.start()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
REPS = 40
e0.record()
for _ in range(REPS):
    lib.sglk_fp8_blockwise_scaled_mm_ex(a.data_ptr(), bt.data_ptr(), sa.data_ptr(), sb.data_ptr(), out.data_ptr(), M, N, K, K, K, N,
                                        sa.stride(0), sa.stride(1), sb.stride(0), sb.stride(1), 0, VAR, NSM, ws.data_ptr(), ws.numel(),
                                        torch.cuda.current_stream().cuda_stream)
e1.record(); e1.synchronize()
_stop = True; th.join()
ms = e0.elapsed_time(e1) / REPS
samples = samples[len(samples) // 3:]
if samples:
    samples.sort()
    print(json.dumps({"ms_per_launch": round(ms, 4), "TFLOPS": round(2.0 * M * N * K / ms / 1e9, 1),
                      "nvml_sm_mhz_median": samples[len(samples) // 2][0], "nvml_power_w_median": sorted(x[1] for x in samples)[len(samples) // 2]}))
d = ws[1 << 20:(1 << 20) + 64 * 8 * 8].view(torch.int64).view(64, 8).cpu()
t0 = int(d[8, 0])
VERBOSE = os.environ.get("VERBOSE", "0") == "1"
if VERBOSE and VAR in (15, 16, 18, 19):
    print("kb: clocks relative to kb 8's MMA issue; mma = issuing thread of CTA 0, epi = epilogue warp 0 of CTA 0")
    for kb in range(8, 32):
        r = [int(v) - t0 for v in d[kb, :8]]
        nxt = [int(v) - t0 for v in d[kb + 1, :8]]
        print(f"{kb:3d}: mma[issue {r[0]:6d} | commit seen {r[2]:6d} (+{r[2]-r[0]:4d})]  epi[first_chunk {r[3]:6d} (+{r[3]-r[2]:4d} after commit) | "
              f"3 chunks {r[4]-r[3]:4d} | release {r[5]-r[4]:4d} | probe+wait+ld {r[7]-r[5]:4d} | last fma {r[6]-r[7]:4d} | "
              f"to next first chunk {nxt[3]-r[6]:4d}]  period {nxt[3]-r[3]:4d}  release->next issue of this stage {int(d[kb + 2, 0]) - t0 - r[5]:4d}")

import statistics
per_epi = [int(d[kb + 1, 3]) - int(d[kb, 3]) for kb in range(8, 60)]
per_mma = [int(d[kb + 1, 0]) - int(d[kb, 0]) for kb in range(8, 60)]
commit = [int(d[kb + 1, 2]) - int(d[kb, 2]) for kb in range(8, 60)]
print("commit-to-commit clocks per K block (as seen by a spinning thread):", commit)
print("tempty ok -> MMAs and commits issued:", [int(d[kb, 1]) - int(d[kb, 0]) for kb in range(8, 60)])
print("issue -> commit seen:", [int(d[kb, 2]) - int(d[kb, 0]) for kb in range(8, 60)])
print(json.dumps({"median_commit_to_commit": statistics.median(commit), "mean_commit_to_commit": sum(commit) / len(commit)}))
blocks_per_worker = -(-((M // 256) * (N // 256)) // ((NSM or 148) // 2)) * (K // 128)
mean_clk = sum(per_mma) / len(per_mma)
print(json.dumps({"implied_sm_mhz": round(mean_clk * blocks_per_worker / (ms * 1e3), 1)}))
print(json.dumps({"variant": VAR, "median_clk_per_kblock_epilogue": statistics.median(per_epi), "median_clk_per_kblock_mma": statistics.median(per_mma),
                  "mean_clk_per_kblock": sum(per_mma) / len(per_mma), "mma_exec_clk": 512}))

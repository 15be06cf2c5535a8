#!/usr/bin/env python
"""bench.py — headline benchmark of the FP8 block-quantized linear hot path on B200.

Workload (BASELINE.json configs[4], the largest config and the one that shards):
    DeepSeek-V3 dense MLP prefill, M=16384 tokens, N=18432, K=7168, bf16 activations,
    e4m3 weights with 128x128 fp32 block scales.
A *step* is one pass of the hot path over one batch:
    per_token_group_quant_8bit (bf16 -> e4m3 + 1x128 fp32 scales, row-major [M, K/128])  ->  fp8_blockwise_scaled_mm
With --gpus N the weight is column-(N-)sharded over the N ranks (one process per GPU, torchrun),
activations replicated; total work is fixed, so scaling is "strong".  The all-gather that
reassembles [M, N] is timed separately ("with_all_gather") because only some consumers need it.

Prints ONE JSON line (rank 0).  `--impl reference` times the reference's own CPU implementation
(the torch oracle restated in oracle/torch_oracle.py; the reference's native code is Intel SYCL and
cannot run here) on a bounded sample of the same workload on ALL host cores (torchrun exports
OMP_NUM_THREADS=1; the CPU legs set the thread count back to os.cpu_count() so that the reference arm is the
same machine at every N).

At N=1 the line also carries `configs`: one sub-record per remaining BASELINE.json config (2: group-quant sweep
points, 3: DSv3 decode projections timed GPU-side in a CUDA graph, 4: Llama-3-70B gate/up prefill), each with its
own roofline.  At N>1 `with_all_gather.fused*` carry `matches_nccl` (bitwise equality with the NCCL-gathered result,
MIN over ranks); a false flag makes the run exit non-zero.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")]

M_FULL, N_FULL, K_FULL = 16384, 18432, 7168
WORKLOAD = "dsv3_dense_mlp_prefill_M16384_N18432_K7168_group128"
METRIC = "FP8 blockwise GEMM TFLOPS (% dense-FP8 peak); group-quant GB/s (% HBM)"
PROFILES = os.path.join(ROOT, "profiles")
_UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def profile_traffic(kernel, want=None):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch, read from the newest committed
    `ncu --set full` summary profiles/rNN*_<kernel>_ncu_metrics.csv (tools/summarize_profiles.py writes them from
    the capture of THIS command).  `want`: substring the captured kernel's name must contain (the template
    arguments of the variant that is timed) — a capture of another variant is not reported as this one's traffic.
    Returns (bytes or None, source description)."""
    import csv
    import glob

    files = sorted(glob.glob(os.path.join(PROFILES, f"r*_{kernel}_ncu_metrics.csv")))
    for path in reversed(files):
        rows = {r[0]: r for r in csv.reader(open(path)) if r}
        name = rows.get("Kernel Name", ["", "", ""])[2]
        if want and want not in name:
            continue
        try:
            rd, wr = rows["dram__bytes_read.sum"], rows["dram__bytes_write.sum"]
            total = float(rd[2]) * _UNIT[rd[1]] + float(wr[2]) * _UNIT[wr[1]]
        except (KeyError, ValueError, IndexError):
            continue
        return int(total), f"{os.path.basename(path)}: {name[:120]}"
    return None, None


# template arguments of fp8_blockwise_gemm_kernel per product variant id (csrc/gemm.cu launch_by_id): what the
# kernel name in an ncu capture must contain for its traffic to be reported against that variant
GEMM_TEMPLATE_ARGS = {13: "<256, 6, 2, __nv_bfloat16, 1, 0, 1, 0, 0, 0, 0", 29: "<256, 5, 2, __nv_bfloat16, 1, 0, 1, 0, 0, 0, 3",
                      0: "<128, 6, 1, __nv_bfloat16, 1, 0, 0, 0, 0, 0, 0"}   # (prefixes: later template parameters have defaults)
FP8_NOMINAL_TFLOPS = 4500.0
HBM_NOMINAL_GBS = 8000.0


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": float(p["hbm_gbs"]), "bf16_tflops": float(p["bf16_tflops"]),
                "bf16_tflops_sustained": float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), "source": "measured"}
    # fallback stated in /opt/skills/guides/B200_PROFILING.md
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU during the timed region (NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.001)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU legs run on rank 0 alone and take the whole host."""
    import torch

    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        pass
    if torch.get_num_threads() != n:
        torch.set_num_threads(n)
    return n


def cpu_oracle_sample(rows, reps_budget_s=12.0, min_reps=2, max_reps=16):
    """Time the CPU oracle (quant + dequantised matmul, all host threads) on the first `rows` tokens of
    the workload.  Returns (best seconds, reps, threads)."""
    import torch

    from oracle import torch_oracle as O

    use_all_host_threads()
    g = torch.Generator().manual_seed(42)
    x = torch.randn(rows, K_FULL, generator=g).to(torch.bfloat16)
    wq_t = ((torch.rand(N_FULL, K_FULL, generator=g) - 0.5) * 896).to(torch.float8_e4m3fn).t()
    ws_t = (torch.rand(K_FULL // 128, N_FULL // 128, generator=g) + 0.5) * 1e-3
    O.fp8_linear_ref(x[:8], wq_t[:, :256], ws_t[:, :2])  # warm the thread pool
    best, reps, t_start = 1e30, 0, time.perf_counter()
    while reps < min_reps or (time.perf_counter() - t_start < reps_budget_s and reps < max_reps):
        t0 = time.perf_counter()
        O.fp8_linear_ref(x, wq_t, ws_t)
        best = min(best, time.perf_counter() - t0)
        reps += 1
    return best, reps, torch.get_num_threads()


def run_reference(args, rank, world):
    """The reference arm: the reference's CPU path (torch oracle) on the host cores, rank 0 only."""
    if rank != 0:
        return
    import torch

    rows = args.cpu_rows
    times = []
    for i in range(args.warmup + args.steps):
        t, _, threads = cpu_oracle_sample(rows, reps_budget_s=0.0, min_reps=1, max_reps=1)
        if i >= args.warmup:
            times.append(t)
    ms = 1e3 * sum(times) / len(times)
    tflops = 2.0 * rows * N_FULL * K_FULL / (ms * 1e-3) / 1e12
    sample = f"first {rows} of {M_FULL} tokens, full N={N_FULL} K={K_FULL}: oracle quant + fp32 dequant torch.mm"
    line = {
        "impl": "reference", "metric": METRIC, "value": tflops, "unit": "TFLOPS", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "fp8_e4m3 (fp32 dequantised on CPU)", "data": "synthetic",
        "config": {"workload": WORKLOAD, "M": M_FULL, "N": N_FULL, "K": K_FULL, "group": 128,
                   "sample_rows_per_step": rows, "parallelism": "cpu threads"},
        "cpu_baseline": {"value": tflops, "unit": "TFLOPS", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": tflops, "unit": "TFLOPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "host_cpus": os.cpu_count(), "torch_threads": torch.get_num_threads(),
    }
    print(json.dumps(line), flush=True)


def other_configs(torch, sgl_kernel, dev, peaks, flush):
    """Sub-records for BASELINE.json configs 2-4 (the headline line is config 5), rank 0 / one GPU.
    Shapes: the reference's own sweeps, benchmark/bench_per_token_group_quant_8bit.py:213-252 (byte model
    3*M*K + 4*M*K/128) and benchmark/bench_fp8_blockwise_gemm.py:16-52 (DeepSeek-V3 / Llama weight shapes).
      * quant, K=7168, M in {4096, 32768}, row- and column-major scales: CUDA events, 256 MiB read-only L2 flush
        between launches, median of 9; roofline = algorithmic bytes / t against the measured copy bandwidth;
      * decode GEMM, N in {2112, 24576}, K=7168, M in {1, 16, 64, 128, 256}: GPU-side time of 16-17 launches in one
        CUDA graph (median of 9 replays); the launches rotate over enough weight copies to exceed 2 x L2
        (17 x 15 MB, 2 x 176 MB) so that every launch streams its weight from HBM; roofline = algorithmic bytes
        M*K + N*K + 2*M*N + scales over t against the measured copy bandwidth; each result is compared with the
        128-row tile kernel (variant 0) on the same inputs;
      * Llama-3-70B gate/up prefill M=8192 N=28672 K=8192: CUDA events with L2 flush, median of 7; roofline tensor."""
    from sgl_kernel import _ops

    lib = sgl_kernel._lib.load()
    hbm, fp8_peak = peaks["hbm_gbs"], 2.0 * peaks["bf16_tflops"]
    recs = []

    def median_events(fn, reps, warm=3):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(reps):
            flush.view(torch.int64).sum()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        return sorted(ts)[len(ts) // 2] * 1e3                      # us

    # ---- config 2: group quant sweep points ----
    K = 7168
    for m in (4096, 32768):
        x = torch.randn(m, K, device=dev, dtype=torch.bfloat16)
        q = torch.empty(m, K, dtype=torch.float8_e4m3fn, device=dev)
        for col in (False, True):
            sc = sgl_kernel.create_per_token_group_quant_fp8_output_scale((m, K), dev, 128, col, col, False)
            us = median_events(lambda: sgl_kernel.sgl_per_token_group_quant_8bit(x, q, sc, 128, 1e-10, -448.0, 448.0,
                                                                                  enable_v2=False), 9)
            by = 3 * m * K + 4 * m * (K // 128)
            recs.append({"config": f"2 quant K=7168 M={m} {'column' if col else 'row'}-major scales", "op": "sgl_per_token_group_quant_8bit",
                         "M": m, "K": K, "us": round(us, 2), "value": by / us / 1e3, "unit": "GB/s",
                         "timing": "events+L2flush, med9",
                         "roofline": {"bound": "hbm", "achieved": by / us / 1e3, "peak": hbm, "unit": "GB/s",
                                      "frac": by / us / 1e3 / hbm, "algorithmic_bytes": by}})
        del x, q

    # ---- config 3: DeepSeek-V3 decode projections, GPU-side in a CUDA graph ----
    def graph_us(fns, reps=9):
        # warm up ON the capture stream: the stream-K workspace is cached per (device, stream) and never created
        # inside a capture, where the op would fall back to whole tiles (sgl_kernel/_ops.py _get_workspace)
        g, st = torch.cuda.CUDAGraph(), torch.cuda.Stream()
        with torch.cuda.stream(st):
            for f in fns[:2]:
                f()
        torch.cuda.synchronize()
        with torch.cuda.stream(st):
            with torch.cuda.graph(g, stream=st):
                for f in fns:
                    f()
        g.replay()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.replay()
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1) / len(fns))
        return sorted(ts)[len(ts) // 2] * 1e3

    for N in (2112, 24576):
        copies = -(-(256 << 20) // (N * K))
        launches = copies * -(-16 // copies)
        bts = [(torch.rand(N, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn) for _ in range(copies)]
        sb = (torch.randn(-(N // -128), K // 128, device=dev) * 1e-3).t()
        for M in (1, 16, 64, 128, 256):
            a = (torch.rand(M, K, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
            sa = torch.randn(M, K // 128, device=dev) * 1e-3
            outs = [torch.empty(M, N, dtype=torch.bfloat16, device=dev) for _ in range(2)]
            y = sgl_kernel.fp8_blockwise_scaled_mm(a, bts[0].t(), sa, sb, torch.bfloat16)
            y0 = _ops._gemm(a, bts[0].t(), sa, sb, torch.bfloat16, 0)
            rel = float((y.float() - y0.float()).norm() / y0.float().norm())
            fns = [(lambda i=i: sgl_kernel.fp8_blockwise_scaled_mm(a, bts[i % copies].t(), sa, sb, torch.bfloat16))
                   for i in range(launches)]
            us = graph_us(fns)
            by = M * K + N * K + 2 * M * N + 4 * (M * (K // 128) + (K // 128) * -(N // -128))
            vid = lib.sglk_gemm_pick_variant(M, N, K)
            recs.append({"config": f"3 DSv3 decode M={M} N={N} K={K}", "op": "fp8_blockwise_scaled_mm", "M": M, "N": N, "K": K,
                         "us": round(us, 2), "value": 2.0 * M * N * K / us / 1e6, "unit": "TFLOPS",
                         "gemm_variant": lib.sglk_gemm_variant_name(vid).decode(), "gemm_variant_id": vid,
                         "rel_err_vs_tile_kernel": rel, "matches_tile_kernel": rel <= 2e-3,
                         "timing": f"graph of {launches} launches over {copies} weight copies (> 2 x L2), med9",
                         "roofline": {"bound": "hbm", "achieved": by / us / 1e3, "peak": hbm, "unit": "GB/s",
                                      "frac": by / us / 1e3 / hbm, "algorithmic_bytes": by}})
            del outs
        del bts

    # ---- config 4: Llama-3-70B gate/up prefill ----
    M, N, K4 = 8192, 28672, 8192
    a = (torch.rand(M, K4, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
    bt = (torch.rand(N, K4, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
    sa = (torch.rand(M, K4 // 128, device=dev) + 0.5) * 1e-3
    sb = ((torch.rand(N // 128, K4 // 128, device=dev) + 0.5) * 1e-3).t()
    us = median_events(lambda: sgl_kernel.fp8_blockwise_scaled_mm(a, bt.t(), sa, sb, torch.bfloat16), 7)
    fl = 2.0 * M * N * K4
    vid = lib.sglk_gemm_pick_variant(M, N, K4)
    recs.append({"config": f"4 Llama-3-70B gate/up prefill M={M} N={N} K={K4}", "op": "fp8_blockwise_scaled_mm", "M": M, "N": N,
                 "K": K4, "us": round(us, 1), "value": fl / us / 1e6, "unit": "TFLOPS",
                 "gemm_variant": lib.sglk_gemm_variant_name(vid).decode(), "gemm_variant_id": vid,
                 "timing": "events+L2flush, med7",
                 "roofline": {"bound": "tensor", "achieved": fl / us / 1e6, "peak": fp8_peak, "unit": "TFLOP/s",
                              "frac": fl / us / 1e6 / fp8_peak, "algorithmic_flops": fl}})
    # ---- config 1: the reference's own CPU-runnable case (tests/test_fp8_blockwise_gemm.py:23-63 torch chain) at M=128
    # N=4096 K=7168: oracle quant + dequantised matmul on the host cores next to quant + GEMM on the GPU, same inputs ----
    try:
        from oracle import torch_oracle as O

        M1, N1, K1 = 128, 4096, 7168
        g = torch.Generator().manual_seed(1)
        x1 = torch.randn(M1, K1, generator=g).to(torch.bfloat16)
        w1 = ((torch.rand(N1, K1, generator=g) - 0.5) * 896).to(torch.float8_e4m3fn)
        s1 = (torch.rand(N1 // 128, K1 // 128, generator=g) + 0.5) * 1e-3
        threads = use_all_host_threads()
        O.fp8_linear_ref(x1[:8], w1.t()[:, :256], s1.t()[:, :2])
        cpu_best = 1e30
        for _ in range(5):
            t0 = time.perf_counter()
            ref1 = O.fp8_linear_ref(x1, w1.t(), s1.t())
            cpu_best = min(cpu_best, time.perf_counter() - t0)
        lin1 = sgl_kernel.Fp8BlockLinear(w1.to(dev), s1.to(dev))
        xd = x1.to(dev)
        us = median_events(lambda: lin1(xd), 9)
        got = lin1(xd).float().cpu()
        rel = float((got - ref1.float()).norm() / ref1.float().norm())
        fl = 2.0 * M1 * N1 * K1
        recs.append({"config": f"1 reference torch path M={M1} N={N1} K={K1}", "op": "quant + fp8_blockwise_scaled_mm",
                     "us": round(us, 2), "value": fl / us / 1e6, "unit": "TFLOPS", "cpu_ms": round(cpu_best * 1e3, 3),
                     "cpu_tflops": fl / cpu_best / 1e12, "cpu_threads": threads, "cpu_kind": "port (oracle torch chain, best of 5)",
                     "rel_err_vs_cpu": rel, "within_tolerance_1e-2": rel <= 1e-2})
    except Exception as e:  # noqa: BLE001
        recs.append({"config": "1 reference torch path", "error": str(e)[:200]})

    # ---- same-run library ceiling: cuBLASLt per-tensor-scaled fp8 (torch._scaled_mm), no block scales, no promotion —
    # the best dense fp8 kernel on this box, beside the assumed 2 x bf16 figure ----
    try:
        one = torch.ones((), device=dev, dtype=torch.float32)
        for (Mc, Nc, Kc) in ((8192, 8192, 8192), (M_FULL, N_FULL, K_FULL), (8192, 28672, 8192)):
            ac = (torch.rand(Mc, Kc, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
            bc = (torch.rand(Nc, Kc, device=dev) - 0.5).mul_(896).to(torch.float8_e4m3fn)
            us = median_events(lambda: torch._scaled_mm(ac, bc.t(), scale_a=one, scale_b=one, out_dtype=torch.bfloat16), 7)
            recs.append({"config": f"library ceiling M={Mc} N={Nc} K={Kc}", "op": "torch._scaled_mm (cuBLASLt, per-tensor fp8 scales)",
                         "M": Mc, "N": Nc, "K": Kc, "us": round(us, 1), "value": 2.0 * Mc * Nc * Kc / us / 1e6, "unit": "TFLOPS"})
            del ac, bc
    except Exception as e:  # noqa: BLE001
        recs.append({"config": "library ceiling", "error": str(e)[:200]})

    # ---- §8 f4: grouped (MoE) blockwise GEMM, DeepSeek-V3 expert shapes, uneven routing (tools/grouped_bench.py) ----
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import grouped_bench

        grouped_bench.Es, grouped_bench.avg = [8, 32], 512
        for r in grouped_bench.records():
            fl = 2.0 * r["sum_m"] * r["N"] * r["K"]
            recs.append({"config": f"f4 grouped MoE GEMM E={r['experts']} {r['proj']} sum_m={r['sum_m']} N={r['N']} K={r['K']}",
                         "op": "fp8_blockwise_scaled_grouped_mm", "us": round(r["ms"] * 1e3, 1), "value": r["tflops"],
                         "unit": "TFLOPS", "dense_op_same_flops_tflops": r["dense_tflops"],
                         "expert0_matches_dense_op": r["expert0_matches_dense_op"],
                         "timing": "events+L2flush, med9",
                         "roofline": {"bound": "tensor", "achieved": r["tflops"], "peak": fp8_peak, "unit": "TFLOP/s",
                                      "frac": r["tflops"] / fp8_peak, "algorithmic_flops": fl}})
    except Exception as e:  # noqa: BLE001
        recs.append({"config": "f4 grouped MoE GEMM", "error": str(e)[:200]})
    return recs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-rows", type=int, default=1024, help="tokens in the CPU-oracle sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the sub-records of BASELINE configs 2-4 (N=1 only)")
    ap.add_argument("--m", type=int, default=M_FULL, help="debug only: override token count")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist

    import sgl_kernel
    from sgl_kernel.sharded import ColumnShardedFp8Linear, shard_bounds

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    M, N, K = args.m, N_FULL, K_FULL
    lo, hi = shard_bounds(N, world, rank)
    n_loc = hi - lo

    # synthetic data, resident in HBM before the timed region
    g = torch.Generator(device=dev).manual_seed(1234)          # identical activations on every rank
    x = torch.randn(M, K, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    gw = torch.Generator(device=dev).manual_seed(77 + rank)
    wq = ((torch.rand(n_loc, K, generator=gw, device=dev) - 0.5) * 896).to(torch.float8_e4m3fn)
    ws = (torch.rand(-(n_loc // -128), K // 128, generator=gw, device=dev) + 0.5) * 1e-3
    layer = sgl_kernel.Fp8BlockLinear(wq, ws)
    sharded = ColumnShardedFp8Linear(layer, N) if world > 1 else None
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def flush_l2():
        # read-only sweep: evicts the step's tensors and leaves CLEAN lines (a memset would leave dirty
        # lines whose write-back competes with the next step's HBM traffic)
        flush.view(torch.int64).sum()


    xq = torch.empty(M, K, dtype=torch.float8_e4m3fn, device=dev)
    xs = sgl_kernel.create_per_token_group_quant_fp8_output_scale((M, K), dev, 128, False, False, False)

    def step():
        sgl_kernel.sgl_per_token_group_quant_8bit(x, xq, xs, 128, 1e-10, -448.0, 448.0, enable_v2=False)
        return layer.forward_quantized(xq, xs)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        flush_l2()
        out = step()
    barrier()

    sampler = ClockSampler(local_rank)
    sampler.start()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    barrier()
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        flush_l2()                                            # L2 flush, outside the per-step events
        ev[i][0].record()
        sgl_kernel.sgl_per_token_group_quant_8bit(x, xq, xs, 128, 1e-10, -448.0, 448.0, enable_v2=False)
        ev[i][1].record()
        out = layer.forward_quantized(xq, xs)
        ev[i][2].record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop()
    t_quant = sum(e[0].elapsed_time(e[1]) for e in ev) / args.steps     # ms
    t_gemm = sum(e[1].elapsed_time(e[2]) for e in ev) / args.steps
    t_step = sum(e[0].elapsed_time(e[2]) for e in ev) / args.steps
    tt = torch.tensor([t_step, t_quant, t_gemm], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_step, t_quant, t_gemm = (float(v) for v in tt.tolist())

    # ---- step + all-gather of the output panels (only for consumers that need [M, N]) ----
    with_gather = None
    gather_mismatch = False

    def same_on_all_ranks(a, b):
        """bitwise equality of two [M, N] results, MIN over ranks"""
        ok = torch.tensor([1 if torch.equal(a, b) else 0], device=dev, dtype=torch.int32)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        return bool(ok.item())

    if world > 1:
        full = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        staged_unused = None
        for _ in range(2):
            sharded.gather(step(), full)
        barrier()
        evg = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for i in range(args.steps):
            flush_l2()
            evg[i][0].record()
            sharded.gather(step(), full)
            evg[i][1].record()
        barrier()
        tg = torch.tensor([sum(a.elapsed_time(b) for a, b in evg) / args.steps], device=dev, dtype=torch.float64)
        dist.all_reduce(tg, op=dist.ReduceOp.MAX)
        with_gather = {"ms_per_step": float(tg.item()), "value": 2.0 * M * N * K / (float(tg.item()) * 1e-3) / 1e12,
                       "unit": "TFLOPS", "collective": "nccl all_gather_into_tensor + panel interleave copy"}
        sharded.gather(step(), full)          # the NCCL-gathered [M, N] the fused forms must reproduce bit for bit
        del staged_unused
        # fused form: the GEMM epilogue TMA-stores each output box into every peer's [M, N] buffer over NVLink
        try:
            sharded.enable_fused_gather(M)
            for _ in range(3):
                sharded.forward_fused_gather(x)
            barrier()
            evf = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
            for i in range(args.steps):
                flush_l2()
                evf[i][0].record()
                sharded.forward_fused_gather(x)
                evf[i][1].record()
            barrier()
            tf = torch.tensor([sum(a.elapsed_time(b) for a, b in evf) / args.steps], device=dev, dtype=torch.float64)
            dist.all_reduce(tf, op=dist.ReduceOp.MAX)
            ok_uni = same_on_all_ranks(sharded.forward_fused_gather(x), full)
            gather_mismatch |= not ok_uni
            with_gather["fused"] = {
                "matches_nccl": ok_uni, "ms_per_step": float(tf.item()), "value": 2.0 * M * N * K / (float(tf.item()) * 1e-3) / 1e12, "unit": "TFLOPS",
                "collective": "none (NCCL-free): GEMM epilogue TMA-stores to peer-mapped symmetric memory + 2 device barriers",
                "nvlink_bytes_in_per_rank": int(M * (N - n_loc) * 2)}
            # NVLS multicast form: each rank sends its panel once, the NVSwitch replicates it
            for _ in range(3):
                sharded.forward_fused_gather(x, multicast=True)
            barrier()
            evm = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
            for i in range(args.steps):
                flush_l2()
                evm[i][0].record()
                sharded.forward_fused_gather(x, multicast=True)
                evm[i][1].record()
            barrier()
            tm = torch.tensor([sum(a.elapsed_time(b) for a, b in evm) / args.steps], device=dev, dtype=torch.float64)
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            ok_mc = same_on_all_ranks(sharded.forward_fused_gather(x, multicast=True), full)
            gather_mismatch |= not ok_mc
            with_gather["fused_multicast"] = {
                "matches_nccl": ok_mc, "ms_per_step": float(tm.item()), "value": 2.0 * M * N * K / (float(tm.item()) * 1e-3) / 1e12, "unit": "TFLOPS",
                "collective": "none (NCCL-free): GEMM epilogue multimem.st to the NVLS multicast mapping + 2 device barriers"}
        except Exception as e:  # noqa: BLE001
            with_gather["fused_error"] = str(e)[:200]
        del full

    # ---- row-sharded activation quantisation + gated GEMM (sgl_kernel/sharded.py forward_sharded_quant): each rank
    # quantises M/g rows, the copy engines push them to the peers behind the GEMM, which starts on its own rows at once.
    # Reported next to `value` (whose step quantises everything on every rank); bitwise equality with that step's
    # output is checked on every rank.
    sharded_quant = None
    if world > 1:
        try:
            if sharded.sharded_quant_supported(M):
                sharded.enable_sharded_quant(M)
                sharded.gated_wait_polls = 16_000_000   # a rank that never delivers fails the bench after ~20-30 s instead of hanging the box
                ref_panel = step()
                ok_sq = True
                for _ in range(3):
                    ok_sq = ok_sq and bool(torch.equal(sharded.forward_sharded_quant(x), ref_panel))
                flag = torch.tensor([1 if ok_sq else 0], device=dev, dtype=torch.int32)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                ok_sq = bool(flag.item())
                gather_mismatch |= not ok_sq
                barrier()
                evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
                for i in range(args.steps):
                    flush_l2()
                    evs[i][0].record()
                    sharded.forward_sharded_quant(x)
                    evs[i][1].record()
                barrier()
                tsq = torch.tensor([sum(a.elapsed_time(b) for a, b in evs) / args.steps], device=dev, dtype=torch.float64)
                dist.all_reduce(tsq, op=dist.ReduceOp.MAX)
                sharded_quant = {"matches_replicated": ok_sq, "ms_per_step": float(tsq.item()),
                                 "value": 2.0 * M * N * K / (float(tsq.item()) * 1e-3) / 1e12, "unit": "TFLOPS",
                                 "transport": "copy engines, ring schedule, per-256-row arrival counters; gated GEMM",
                                 "nvlink_bytes_in_per_rank": int((M - M // world) * (K + (K // 128) * 4))}
        except Exception as e:  # noqa: BLE001
            sharded_quant = {"error": str(e)[:200]}

    # ---- end to end through the public API with HOST buffers (pinned), copies inside the timed region ----
    e2e = None
    if not args.no_e2e:
        x_host = torch.empty(M, K, dtype=torch.bfloat16).pin_memory()
        x_host.copy_(x)
        out_hosts = [torch.empty(M, n_loc, dtype=torch.bfloat16).pin_memory() for _ in range(2)]
        # public host-buffer API: H2D / quant+GEMM / D2H overlapped; N>1: every rank uploads M/g rows, quantises them and
        # all-gathers codes + scales over NVLink instead of pushing g copies of x through the host's PCIe root
        runner = (sgl_kernel.HostStreamedShardedFp8Linear(sharded, M) if world > 1
                  else sgl_kernel.HostStreamedFp8Linear(layer, M))
        e2e_steps = max(4, min(args.steps, 12))
        for i in range(2):
            runner.submit(x_host, out_hosts[i & 1])
        runner.synchronize()
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            runner.submit(x_host, out_hosts[i & 1])
        runner.synchronize()
        te_local = (time.perf_counter() - t0) * 1e3 / e2e_steps
        barrier()
        te = torch.tensor([te_local], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        # sanity: the host result of the last step is the device result
        e2e_ok = bool(torch.equal(out_hosts[(e2e_steps - 1) & 1][:64].to(dev), out[:64]))
        # denominator: the same host->device and device->host byte volume per step as plain asynchronous copies
        # (one cudaMemcpyAsync per direction per step, through Tensor.copy_(non_blocking=True)) on the same pinned
        # buffers, two streams so that both PCIe directions run concurrently, every rank at once — no compute,
        # no collective.  frac = copy time / e2e time: 1.0 means the pipeline is at the PCIe roofline.
        chunk = -(-M // world)
        r_lo, r_hi = min(rank * chunk, M), min((rank + 1) * chunk, M)
        x_stage = torch.empty(max(r_hi - r_lo, 1), K, dtype=torch.bfloat16, device=dev)
        s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

        def copies_only():
            with torch.cuda.stream(s_up):
                if r_hi > r_lo:
                    x_stage[: r_hi - r_lo].copy_(x_host[r_lo:r_hi], non_blocking=True)
            with torch.cuda.stream(s_dn):
                out_hosts[0].copy_(out, non_blocking=True)

        for _ in range(2):
            copies_only()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            copies_only()
        s_up.synchronize()
        s_dn.synchronize()
        tc_local = (time.perf_counter() - t0) * 1e3 / e2e_steps
        barrier()
        tcp = torch.tensor([tc_local], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tcp, op=dist.ReduceOp.MAX)
        t_copy = float(tcp.item())
        up_bytes, dn_bytes = int(M * K * 2), int(M * N * 2)
        del x_stage
        e2e = {"value": 2.0 * M * N * K / (float(te.item()) * 1e-3) / 1e12, "unit": "TFLOPS",
               "ms_per_step": float(te.item()), "steps": e2e_steps,
               "h2d_bytes_per_step": int(x_host.numel() * 2), "d2h_bytes_per_step": int(M * N * 2),
               "matches_device_result": e2e_ok,
               "roofline": {"bound": "pcie", "ms_copies_only": t_copy, "frac": t_copy / float(te.item()),
                            "h2d_gbs": up_bytes / t_copy / 1e6, "d2h_gbs": dn_bytes / t_copy / 1e6,
                            "what": "same H2D + D2H bytes per step as plain async copies on the same pinned buffers, both "
                                    "directions concurrently on two streams, all ranks at once, no compute; whole-job GB/s"},
               "frac": t_copy / float(te.item()),
               "note": ("sgl_kernel.HostStreamedFp8Linear: pinned host x -> HBM, quant + GEMM ops, result panel -> pinned host, "
                        "every step; the three phases of consecutive steps overlap on three streams; wall clock over "
                        "all steps incl. final sync; bytes are whole-job (all ranks)") if world == 1 else
                       ("sgl_kernel.HostStreamedShardedFp8Linear: every rank uploads M/g rows of the pinned host x, quantises "
                        "them, NCCL all-gathers codes + scales over NVLink, GEMMs its N/g panel and downloads it to pinned "
                        "host memory, every step; phases of consecutive steps overlap on three streams; wall clock over all "
                        "steps incl. final sync, max over ranks; bytes are whole-job (all ranks)")}

    # CPU baseline on rank 0 with every host thread; the other ranks sleep on the rendezvous store meanwhile (a
    # device-side barrier would keep one host thread per rank spinning in cudaStreamSynchronize and steal cores)
    cpu_sample = None
    store = dist.distributed_c10d._get_default_store() if world > 1 else None
    if rank == 0 and not args.no_cpu_baseline:
        cpu_sample = cpu_oracle_sample(args.cpu_rows)
    if store is not None:
        import datetime

        if rank == 0:
            store.set("sglk_bench_cpu_done", "1")
        else:
            store.wait(["sglk_bench_cpu_done"], datetime.timedelta(minutes=15))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        if gather_mismatch:
            raise SystemExit(1)
        return

    lib = sgl_kernel._lib.load()
    peaks = load_peaks()
    fp8_peak = 2.0 * peaks["bf16_tflops"]
    gemm_tflops_rank = 2.0 * M * n_loc * K / (t_gemm * 1e-3) / 1e12
    quant_bytes = M * K * 3 + M * (K // 128) * 4
    quant_gbs = quant_bytes / (t_quant * 1e-3) / 1e9
    value = 2.0 * M * N * K / (t_step * 1e-3) / 1e12
    # N > 1: two exact implementations of the same step exist (identical inputs, bitwise-identical output panels, both timed
    # over the same K steps above): activations quantised on every rank, or row-sharded and exchanged over NVLink under the
    # GEMM.  `value` is the faster one where the sharded form wins by more than 2 % (4 GPUs on this hardware); both are
    # reported.  gemm_ms / quant_ms / the rooflines always describe the replicated form's kernels.
    value_form = "activations quantised on every rank (no collective)" if world > 1 else "single GPU: quant + GEMM"
    replicated = {"value": value, "ms_per_step": t_step}
    if (sharded_quant and sharded_quant.get("matches_replicated") and
            sharded_quant.get("ms_per_step", 1e30) < 0.98 * t_step):
        value_form = "row-sharded quantisation exchanged over NVLink (copy engines) under the gated GEMM"
        t_step = sharded_quant["ms_per_step"]
        value = sharded_quant["value"]

    # ncu DRAM traffic of the kernels that were just timed (only meaningful for the full single-GPU shape the
    # capture was taken on; the name check ties it to the variant's template instantiation)
    gemm_traffic = gemm_traffic_src = quant_traffic = quant_traffic_src = None
    variant_id = lib.sglk_gemm_pick_variant(M, n_loc, K)
    if M == M_FULL:
        quant_traffic, quant_traffic_src = profile_traffic("quant", "ptgq_kernel")
        if world == 1:
            gemm_traffic, gemm_traffic_src = profile_traffic("gemm", GEMM_TEMPLATE_ARGS.get(variant_id, "\0"))

    cpu_baseline = None
    if cpu_sample is not None:
        best, reps, threads = cpu_sample
        cpu_baseline = {
            "value": 2.0 * args.cpu_rows * N * K / best / 1e12, "unit": "TFLOPS", "cores": threads, "kind": "port",
            "sample": f"first {args.cpu_rows} of {M} tokens, full N={N} K={K}: oracle quant + fp32 dequant torch.mm, "
                      f"best of {reps} ({best:.2f} s each)", "host_cpus": os.cpu_count(),
        }

    line = {
        "metric": METRIC, "value": value, "unit": "TFLOPS", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "fp8_e4m3 x fp8_e4m3 -> fp32 accumulate -> bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "M": M, "N": N, "K": K, "group": 128, "n_per_rank": n_loc,
                   "parallelism": f"column(N)-sharded x{world}, activations replicated" if world > 1 else "single GPU",
                   "step": "per_token_group_quant_8bit (row-major scales) + fp8_blockwise_scaled_mm; output stays N-sharded",
                   "step_form": value_form,
                   "l2": "256 MiB read-only flush between steps (outside the events); inputs 235 MB + weights > L2",
                   "gemm_variant": lib.sglk_gemm_variant_name(variant_id).decode(), "gemm_variant_id": variant_id},
        "gemm_tflops": 2.0 * M * N * K / (t_gemm * 1e-3) / 1e12, "gemm_ms": t_gemm,
        "quant_gbps": quant_gbs, "quant_ms": t_quant,
        "roofline": {"bound": "tensor", "achieved": gemm_tflops_rank, "peak": fp8_peak, "unit": "TFLOP/s",
                     "frac": gemm_tflops_rank / fp8_peak,
                     "traffic": gemm_traffic, "traffic_source": gemm_traffic_src,
                     "kernel": "fp8_blockwise_gemm_kernel", "algorithmic_flops_per_launch": 2.0 * M * n_loc * K,
                     "peak_source": f"2 x {peaks['source']} cuBLAS bf16 burst ({peaks['bf16_tflops']} TF/s): fp8 dense = 2x bf16 rate; "
                                    f"frac of nominal 4500 TF/s = {gemm_tflops_rank / FP8_NOMINAL_TFLOPS:.3f}"},
        "roofline_quant": {"bound": "hbm", "achieved": quant_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                           "frac": quant_gbs / peaks["hbm_gbs"],
                           "traffic": quant_traffic, "traffic_source": quant_traffic_src, "kernel": "ptgq_kernel",
                           "algorithmic_bytes_per_launch": quant_bytes,
                           "peak_source": f"{peaks['source']} copy bandwidth; frac of nominal 8000 GB/s = {quant_gbs / HBM_NOMINAL_GBS:.3f}"},
        "cpu_baseline": cpu_baseline, "e2e": e2e, "with_all_gather": with_gather, "sharded_quant": sharded_quant,
        "replicated_quant": replicated,
        "gpu_launches": 2 * args.steps * world, "clocks": clocks, "wall_s_timed_region": t_wall,
    }
    if world == 1 and not args.no_configs and M == M_FULL:
        del out, x, xq, xs, layer, wq, ws
        torch.cuda.empty_cache()
        t0 = time.perf_counter()
        line["configs"] = other_configs(torch, sgl_kernel, dev, peaks, flush)
        line["configs_wall_s"] = round(time.perf_counter() - t0, 2)
        ceil = [r for r in line["configs"] if r.get("config") == f"library ceiling M={M_FULL} N={N_FULL} K={K_FULL}" and "value" in r]
        if ceil:   # the headline GEMM against the best library fp8 kernel measured in this same run
            line["roofline"]["library_ceiling"] = {"tflops": ceil[0]["value"], "op": ceil[0]["op"],
                                                   "frac_of_library": line["gemm_tflops"] / ceil[0]["value"]}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if gather_mismatch:
        raise SystemExit("a fused multi-GPU form differs from its reference (see with_all_gather.*.matches_nccl, sharded_quant.matches_replicated)")


if __name__ == "__main__":
    main()

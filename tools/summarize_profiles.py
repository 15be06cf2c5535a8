#!/usr/bin/env python
"""Turn the ncu artefacts in gpurun_out/ into the small tracked summaries under profiles/.
usage: summarize_profiles.py r01"""
import csv
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R = sys.argv[1] if len(sys.argv) > 1 else "r01"
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")
os.makedirs(P, exist_ok=True)

KEEP = ("gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__cluster_size", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__ops_path_tensor_src_fp4_fp6_fp8_dst_fp32.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed")


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    return rows[0], rows[1], rows[2:]


def dump_kernel(name):
    rep = os.path.join(G, f"{name}_{R}.ncu-rep")
    if not os.path.exists(rep):
        return
    hdr, units, vals = raw(rep)
    with open(os.path.join(P, f"{R}_{name}_ncu_metrics.csv"), "w") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + [f"launch{i}" for i in range(len(vals))])
        for i, h in enumerate(hdr):
            if h in KEEP or "issue_stalled" in h and "per_issue_active" in h or h == "Kernel Name":
                w.writerow([h, units[i]] + [v[i][:160] for v in vals])
    # top stall locations from the source page
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[2:] if len(r) == len(hdr)]
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[ix["# Samples"]] or 0) for r in data) or 1
    with open(os.path.join(P, f"{R}_{name}_ncu_hot_instructions.csv"), "w") as f:
        w = csv.writer(f)
        w.writerow(["address", "samples", "pct", "sass", "top_stall", "second_stall"])
        for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]] or 0))[:40]:
            s = int(r[ix["# Samples"]] or 0)
            st = sorted(((h, int(r[ix[h]] or 0)) for h in stalls), key=lambda kv: -kv[1])[:2]
            w.writerow([r[ix["Address"]][-6:], s, f"{100 * s / tot:.1f}", r[ix["Source"]].strip()[:100],
                        f"{st[0][0]}={st[0][1]}", f"{st[1][0]}={st[1][1]}"])


def dump_launches():
    src = os.path.join(G, f"launches_{R}.csv")
    if not os.path.exists(src):
        return
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    hdr = rows[0]
    ix = {h: i for i, h in enumerate(hdr)}
    with open(os.path.join(P, f"{R}_launches.csv"), "w") as f:
        w = csv.writer(f)
        w.writerow(["id", "kernel", "block", "grid", "duration_us"])
        for r in rows[1:]:
            v = float(r[ix["Metric Value"]])
            if r[ix["Metric Unit"]] in ("ns", "nsecond"):
                v /= 1e3
            w.writerow([r[ix["ID"]], r[ix["Kernel Name"]].split("(")[0][:90], r[ix["Block Size"]], r[ix["Grid Size"]], f"{v:.1f}"])


dump_launches()
dump_kernel("gemm")
dump_kernel("quant")
dump_kernel("decode")
dump_kernel("decoden")
dump_kernel("decode64")
dump_kernel("decode128")
dump_kernel("decode256")
dump_kernel("decoden128")
dump_kernel("grouped")
dump_kernel("gated")
dump_kernel("tokenquant")
print(sorted(os.listdir(P)))
